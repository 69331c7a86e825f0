// host.cu — HOST-buffer entry points: the reference-facing calls (numpy in, numpy out).
//
// VampPreviewPlanner hands VAMP Python lists one configuration at a time
// (TyGrit/planning/fetch/vamp_preview.py:159-163, 306-338).  Here a whole batch arrives in host
// memory; it is cut into chunks that alternate between two streams so the H2D copy of chunk i+1,
// the kernels of chunk i and the D2H copy of chunk i-1 overlap.  With pinned host buffers the
// path is PCIe-bound (44 B in + 1 B out per configuration).
#include <string.h>

#include "common.cuh"

static int tg_host_prepare(tg_env* env) {
  if (env->streams[0]) return TG_OK;
  env->chunk = 1 << 20;  // configurations (or 1/4 as many edges) per chunk
  for (int i = 0; i < 2; ++i) {
    TG_CUDA(cudaStreamCreateWithFlags(&env->streams[i], cudaStreamNonBlocking));
    TG_CUDA(cudaEventCreateWithFlags(&env->events[i], cudaEventDisableTiming));
    TG_CUDA(cudaMalloc(&env->dev_in[i], sizeof(float) * 22 * (size_t)env->chunk));
    TG_CUDA(cudaMalloc(&env->dev_out[i], (size_t)env->chunk + 64));
    TG_CUDA(cudaMalloc(&env->dev_clear[i], sizeof(float) * (size_t)env->chunk));
  }
  TG_CUDA(cudaMalloc(&env->dev_checks, sizeof(int64_t) * TG_HOST_MAX_CHUNKS));
  TG_CUDA(cudaMallocHost(&env->pin_checks, sizeof(int64_t) * TG_HOST_MAX_CHUNKS));
  TG_CUDA(cudaMallocHost(&env->pin_small, sizeof(float) * 22 * TG_HOST_SMALL));
  TG_CUDA(cudaMallocHost(&env->pin_small_out, 5 * TG_HOST_SMALL + 64));
  return TG_OK;
}

extern "C" int tg_check_configs_host(tg_env* env, const float* arm, const float* base, int64_t B, uint8_t* valid, float* clearance) {
  TG_CHECK_ARG(env != nullptr, "tg_check_configs_host: env is NULL");
  TG_CHECK_ARG(B >= 0 && (B == 0 || (arm && base && valid)), "tg_check_configs_host: NULL buffer");
  if (B == 0) return TG_OK;
  TG_CUDA(cudaSetDevice(env->device));
  int rc = tg_host_prepare(env);
  if (rc) return rc;
  TG_CUDA(cudaDeviceSynchronize());  // the world may have been built on another stream
  if (B <= TG_HOST_SMALL) {
    // small batch (the one-configuration-at-a-time calls of the reference surface land here): pack into pinned staging,
    // ONE H2D copy, the kernels, ONE D2H copy — a pageable-memory cudaMemcpyAsync costs more than the kernels
    cudaStream_t st = env->streams[0];
    float* d_arm = env->dev_in[0];
    float* d_base = d_arm + 8 * B;
    memcpy(env->pin_small, arm, sizeof(float) * 8 * B);
    memcpy(env->pin_small + 8 * B, base, sizeof(float) * 3 * B);
    TG_CUDA(cudaMemcpyAsync(d_arm, env->pin_small, sizeof(float) * 11 * B, cudaMemcpyHostToDevice, st));
    rc = tg_check_configs_slot(env, d_arm, d_base, B, env->dev_out[0], clearance ? env->dev_clear[0] : nullptr, st, 0);
    if (rc) return rc;
    TG_CUDA(cudaMemcpyAsync(env->pin_small_out, env->dev_out[0], (size_t)B, cudaMemcpyDeviceToHost, st));
    if (clearance) TG_CUDA(cudaMemcpyAsync(env->pin_small, env->dev_clear[0], sizeof(float) * B, cudaMemcpyDeviceToHost, st));
    TG_CUDA(cudaStreamSynchronize(st));
    memcpy(valid, env->pin_small_out, (size_t)B);
    if (clearance) memcpy(clearance, env->pin_small, sizeof(float) * B);
    return TG_OK;
  }
  const int64_t chunk = env->chunk;
  int c = 0;
  for (int64_t off = 0; off < B; off += chunk, ++c) {
    const int b = c & 1;
    const int64_t n = (B - off) < chunk ? (B - off) : chunk;
    cudaStream_t st = env->streams[b];
    float* d_arm = env->dev_in[b];
    float* d_base = env->dev_in[b] + 8 * chunk;
    TG_CUDA(cudaMemcpyAsync(d_arm, arm + 8 * off, sizeof(float) * 8 * n, cudaMemcpyHostToDevice, st));
    TG_CUDA(cudaMemcpyAsync(d_base, base + 3 * off, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st));
    rc = tg_check_configs_slot(env, d_arm, d_base, n, env->dev_out[b], clearance ? env->dev_clear[b] : nullptr, st, b);
    if (rc) return rc;
    TG_CUDA(cudaMemcpyAsync(valid + off, env->dev_out[b], (size_t)n, cudaMemcpyDeviceToHost, st));
    if (clearance) TG_CUDA(cudaMemcpyAsync(clearance + off, env->dev_clear[b], sizeof(float) * n, cudaMemcpyDeviceToHost, st));
  }
  TG_CUDA(cudaStreamSynchronize(env->streams[0]));
  TG_CUDA(cudaStreamSynchronize(env->streams[1]));
  return TG_OK;
}

extern "C" int tg_check_edges_host(tg_env* env, const float* arm0, const float* base0, const float* arm1, const float* base1,
                                   int64_t E, int rake, float resolution, uint8_t* valid, int64_t* n_checks) {
  TG_CHECK_ARG(env != nullptr, "tg_check_edges_host: env is NULL");
  TG_CHECK_ARG(E >= 0 && (E == 0 || (arm0 && base0 && arm1 && base1 && valid)), "tg_check_edges_host: NULL buffer");
  if (n_checks) *n_checks = 0;
  if (E == 0) return TG_OK;
  TG_CUDA(cudaSetDevice(env->device));
  int rc = tg_host_prepare(env);
  if (rc) return rc;
  TG_CUDA(cudaDeviceSynchronize());
  if (E <= TG_HOST_SMALL) {  // small batch: one packed H2D copy, one D2H copy (verdict bytes + the sample counter)
    cudaStream_t st = env->streams[0];
    float* d_a0 = env->dev_in[0];
    float* d_a1 = d_a0 + 8 * E;
    float* d_b0 = d_a1 + 8 * E;
    float* d_b1 = d_b0 + 3 * E;
    memcpy(env->pin_small, arm0, sizeof(float) * 8 * E);
    memcpy(env->pin_small + 8 * E, arm1, sizeof(float) * 8 * E);
    memcpy(env->pin_small + 16 * E, base0, sizeof(float) * 3 * E);
    memcpy(env->pin_small + 19 * E, base1, sizeof(float) * 3 * E);
    TG_CUDA(cudaMemcpyAsync(d_a0, env->pin_small, sizeof(float) * 22 * E, cudaMemcpyHostToDevice, st));
    const size_t cnt_off = ((size_t)E + 7) & ~(size_t)7;
    int64_t* d_cnt = reinterpret_cast<int64_t*>(env->dev_out[0] + cnt_off);
    rc = tg_check_edges_slot(env, d_a0, d_b0, d_a1, d_b1, E, rake, resolution, env->dev_out[0], d_cnt, st, 0);
    if (rc) return rc;
    TG_CUDA(cudaMemcpyAsync(env->pin_small_out, env->dev_out[0], cnt_off + sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    TG_CUDA(cudaStreamSynchronize(st));
    memcpy(valid, env->pin_small_out, (size_t)E);
    if (n_checks) memcpy(n_checks, env->pin_small_out + cnt_off, sizeof(int64_t));
    return TG_OK;
  }
  const int64_t chunk = env->chunk / 4;
  const int64_t n_chunks = (E + chunk - 1) / chunk;
  // one counter per chunk, summed on the host at the end; preallocated (a cudaMalloc + cudaMallocHost pair per call
  // costs more than a small batch's kernels)
  const bool own = n_chunks > TG_HOST_MAX_CHUNKS;
  int64_t* d_checks = env->dev_checks;
  int64_t* h_checks = env->pin_checks;
  if (own) {
    TG_CUDA(cudaMalloc(&d_checks, sizeof(int64_t) * n_chunks));
    TG_CUDA(cudaMallocHost(&h_checks, sizeof(int64_t) * n_chunks));
  }
  int c = 0;
  for (int64_t off = 0; off < E; off += chunk, ++c) {
    const int b = c & 1;
    const int64_t n = (E - off) < chunk ? (E - off) : chunk;
    cudaStream_t st = env->streams[b];
    float* d_a0 = env->dev_in[b];
    float* d_b0 = d_a0 + 8 * chunk;
    float* d_a1 = d_b0 + 3 * chunk;
    float* d_b1 = d_a1 + 8 * chunk;
    d_b1 = (float*)(((uintptr_t)d_b1 + 15) & ~(uintptr_t)15);
    TG_CUDA(cudaMemcpyAsync(d_a0, arm0 + 8 * off, sizeof(float) * 8 * n, cudaMemcpyHostToDevice, st));
    TG_CUDA(cudaMemcpyAsync(d_b0, base0 + 3 * off, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st));
    TG_CUDA(cudaMemcpyAsync(d_a1, arm1 + 8 * off, sizeof(float) * 8 * n, cudaMemcpyHostToDevice, st));
    TG_CUDA(cudaMemcpyAsync(d_b1, base1 + 3 * off, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st));
    rc = tg_check_edges_slot(env, d_a0, d_b0, d_a1, d_b1, n, rake, resolution, env->dev_out[b], d_checks + c, st, b);
    if (rc) break;
    TG_CUDA(cudaMemcpyAsync(valid + off, env->dev_out[b], (size_t)n, cudaMemcpyDeviceToHost, st));
    TG_CUDA(cudaMemcpyAsync(h_checks + c, d_checks + c, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  }
  cudaStreamSynchronize(env->streams[0]);
  cudaStreamSynchronize(env->streams[1]);
  if (rc == TG_OK && n_checks)
    for (int64_t i = 0; i < n_chunks; ++i) *n_checks += h_checks[i];
  if (own) {
    cudaFree(d_checks);
    cudaFreeHost(h_checks);
  }
  if (rc) return rc;
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}
