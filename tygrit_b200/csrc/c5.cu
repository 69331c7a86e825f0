// c5.cu — the receding-horizon replanning step as ONE CUDA graph, driven and timed from C.
//
// BASELINE config 5 ("per-step cloud rebuild + 1M edge checks per step over 100 scheduler steps") mimics the
// reference's Scheduler loop: every iteration calls planner.update_environment (a full CAPT rebuild behind
// TyGrit/planning/fetch/vamp_preview.py:183-240) and then plans, i.e. validates edges
// (planning/fetch/algorithms.py:37,51).  Here one step = [fresh edges] -> tg_env_build -> tg_check_edges, captured
// once per cloud buffer with fixed workspace bounds (the build then never synchronises) and replayed: no Python and
// no per-launch driver overhead inside the timed loop.  The edges of a step are generated ON THE DEVICE by a counter-
// keyed hash (a device-side step counter is part of the graph), so every replay validates different edges.
#include <math.h>

#include <vector>

#include "common.cuh"
#include "fetch_model.cuh"

__constant__ float TGC5_LO[8] = {0.0f, -1.6056f, -1.221f, -3.14159265f, -2.251f, -3.14159265f, -2.16f, -3.14159265f};
__constant__ float TGC5_HI[8] = {0.38615f, 1.6056f, 1.518f, 3.14159265f, 2.251f, 3.14159265f, 2.16f, 3.14159265f};

__device__ __forceinline__ uint32_t tgc5_hash(uint32_t a, uint32_t b, uint32_t c) {  // three rounds of a 32-bit mixer
  uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du;
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ float tgc5_u01(uint32_t h) { return (float)(h >> 8) * (1.0f / 16777216.0f); }

// arm-only edges at a random base (plan_arm_rrtc, algorithms.py:23): both endpoints uniform in the joint limits
__global__ void __launch_bounds__(256)
k_c5_edges(const unsigned* __restrict__ step, uint32_t seed, float4 box, int E, float* __restrict__ arm0, float* __restrict__ base0,
           float* __restrict__ arm1, float* __restrict__ base1) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const uint32_t s = *step + seed;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    arm0[8 * (size_t)e + k] = TGC5_LO[k] + (TGC5_HI[k] - TGC5_LO[k]) * tgc5_u01(tgc5_hash(s, (uint32_t)e, 2 * k));
    arm1[8 * (size_t)e + k] = TGC5_LO[k] + (TGC5_HI[k] - TGC5_LO[k]) * tgc5_u01(tgc5_hash(s, (uint32_t)e, 2 * k + 1));
  }
  const float bx = box.x + (box.y - box.x) * tgc5_u01(tgc5_hash(s, (uint32_t)e, 16));
  const float by = box.z + (box.w - box.z) * tgc5_u01(tgc5_hash(s, (uint32_t)e, 17));
  const float bt = -3.14159265f + 6.2831853f * tgc5_u01(tgc5_hash(s, (uint32_t)e, 18));
  base0[3 * (size_t)e] = bx; base0[3 * (size_t)e + 1] = by; base0[3 * (size_t)e + 2] = bt;
  base1[3 * (size_t)e] = bx; base1[3 * (size_t)e + 1] = by; base1[3 * (size_t)e + 2] = bt;
}
__global__ void k_c5_bump(unsigned* step) { *step += 1u; }
__global__ void __launch_bounds__(256) k_c5_count_valid(const uint8_t* __restrict__ valid, int E, unsigned long long* out) {
  unsigned n = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) n += valid[i] != 0;
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(out, (unsigned long long)n);
}

#define C5_CUDA(expr)                                                                               \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess) {                                                                        \
      tg_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);     \
      rc = TG_E_CUDA;                                                                               \
      goto done;                                                                                    \
    }                                                                                               \
  } while (0)

// out8: [0] ms per step (graph replay, mean over `steps`), [1] ms per build-only graph replay, [2] edges valid fraction
// (all timed steps), [3] configuration checks per step (mean), [4] graphs captured, [5] kernel + memset nodes per step
// graph, [6] min step ms, [7] max step ms (per-step events every 10th step).
extern "C" int tg_c5_run(tg_env* env, const float* const* clouds, const int64_t* n_points, int n_clouds, int64_t E, int steps,
                         int warmup, uint64_t seed, const float* base_box4, int rake, float resolution, double* out8) {
  TG_CHECK_ARG(env != nullptr && clouds != nullptr && n_points != nullptr && n_clouds >= 1 && n_clouds <= 64 && out8 != nullptr,
               "tg_c5_run: bad argument");
  TG_CHECK_ARG(E >= 1 && E <= (1 << 24) && steps >= 1 && warmup >= 0 && base_box4 != nullptr, "tg_c5_run: bad sizes");
  TG_CHECK_ARG(env->fixed_bounds, "tg_c5_run: fix the workspace box first (tg_env_set_bounds): a build that reads the cloud bounds back cannot be captured");
  int rc = TG_OK;
  cudaStream_t st = nullptr;
  float *arm0 = nullptr, *arm1 = nullptr, *base0 = nullptr, *base1 = nullptr;
  uint8_t* valid = nullptr;
  unsigned* step_ctr = nullptr;
  unsigned long long* acc = nullptr;  // [0] valid edges, [1] config checks (int64 written by tg_check_edges), [2] total checks
  std::vector<cudaGraphExec_t> step_graphs, build_graphs;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  const float4 box = make_float4(base_box4[0], base_box4[1], base_box4[2], base_box4[3]);
  size_t n_nodes = 0;
  double t_min = 1e30, t_max = 0;
  for (int i = 0; i < 8; ++i) out8[i] = 0;
  C5_CUDA(cudaSetDevice(env->device));
  C5_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  C5_CUDA(cudaMalloc(&arm0, sizeof(float) * 8 * E));
  C5_CUDA(cudaMalloc(&arm1, sizeof(float) * 8 * E));
  C5_CUDA(cudaMalloc(&base0, sizeof(float) * 3 * E));
  C5_CUDA(cudaMalloc(&base1, sizeof(float) * 3 * E));
  C5_CUDA(cudaMalloc(&valid, (size_t)E));
  C5_CUDA(cudaMalloc(&step_ctr, sizeof(unsigned)));
  C5_CUDA(cudaMalloc(&acc, 4 * sizeof(unsigned long long)));
  C5_CUDA(cudaMemsetAsync(step_ctr, 0, sizeof(unsigned), st));
  C5_CUDA(cudaMemsetAsync(acc, 0, 4 * sizeof(unsigned long long), st));
  for (int k = 0; k < 4; ++k) C5_CUDA(cudaEventCreate(&ev[k]));
  // one eager pass per cloud: sizes the texture array (allocation is not capturable) and warms the kernels up
  for (int c = 0; c < n_clouds; ++c) {
    rc = tg_env_build(env, clouds[c], n_points[c], st);
    if (rc) goto done;
  }
  C5_CUDA(cudaStreamSynchronize(st));
  for (int c = 0; c < n_clouds; ++c) {
    for (int kind = 0; kind < 2; ++kind) {  // 0: whole step, 1: build only
      cudaGraph_t graph = nullptr;
      cudaGraphExec_t exec = nullptr;
      C5_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      if (kind == 0) {
        k_c5_edges<<<tg_div_up(E, 256), 256, 0, st>>>(step_ctr, (uint32_t)seed, box, (int)E, arm0, base0, arm1, base1);
        k_c5_bump<<<1, 1, 0, st>>>(step_ctr);
      }
      rc = tg_env_build(env, clouds[c], n_points[c], st);
      if (rc == TG_OK && kind == 0) {
        rc = tg_check_edges(env, arm0, base0, arm1, base1, E, rake, resolution, valid, (int64_t*)(acc + 1), st);
        k_c5_count_valid<<<env->sm_count * 4, 256, 0, st>>>(valid, (int)E, acc);
      }
      cudaError_t ce = cudaStreamEndCapture(st, &graph);
      if (rc) {
        if (graph) cudaGraphDestroy(graph);
        goto done;
      }
      C5_CUDA(ce);
      if (kind == 0 && c == 0) cudaGraphGetNodes(graph, nullptr, &n_nodes);
      C5_CUDA(cudaGraphInstantiate(&exec, graph, 0));
      cudaGraphDestroy(graph);
      (kind == 0 ? step_graphs : build_graphs).push_back(exec);
    }
  }
  for (int i = 0; i < warmup; ++i) C5_CUDA(cudaGraphLaunch(step_graphs[i % n_clouds], st));
  C5_CUDA(cudaMemsetAsync(acc, 0, 4 * sizeof(unsigned long long), st));
  C5_CUDA(cudaStreamSynchronize(st));
  {
    unsigned long long checks_total = 0, h_acc[4];
    C5_CUDA(cudaEventRecord(ev[0], st));
    for (int i = 0; i < steps; ++i) {
      const bool probe = (i % 10) == 0;
      if (probe) C5_CUDA(cudaEventRecord(ev[2], st));
      C5_CUDA(cudaGraphLaunch(step_graphs[i % n_clouds], st));
      if (probe) {  // per-step spread (one in ten steps pays a stream sync)
        C5_CUDA(cudaEventRecord(ev[3], st));
        C5_CUDA(cudaEventSynchronize(ev[3]));
        float ms = 0;
        C5_CUDA(cudaEventElapsedTime(&ms, ev[2], ev[3]));
        t_min = ms < t_min ? ms : t_min;
        t_max = ms > t_max ? ms : t_max;
        C5_CUDA(cudaMemcpy(h_acc, acc, sizeof(h_acc), cudaMemcpyDeviceToHost));
        checks_total += h_acc[1];
      }
    }
    C5_CUDA(cudaEventRecord(ev[1], st));
    C5_CUDA(cudaEventSynchronize(ev[1]));
    float ms = 0;
    C5_CUDA(cudaEventElapsedTime(&ms, ev[0], ev[1]));
    out8[0] = ms / steps;
    C5_CUDA(cudaMemcpy(h_acc, acc, sizeof(h_acc), cudaMemcpyDeviceToHost));
    out8[2] = (double)h_acc[0] / ((double)E * steps);
    out8[3] = (double)checks_total / (double)((steps + 9) / 10);
    C5_CUDA(cudaEventRecord(ev[0], st));
    for (int i = 0; i < steps; ++i) C5_CUDA(cudaGraphLaunch(build_graphs[i % n_clouds], st));
    C5_CUDA(cudaEventRecord(ev[1], st));
    C5_CUDA(cudaEventSynchronize(ev[1]));
    C5_CUDA(cudaEventElapsedTime(&ms, ev[0], ev[1]));
    out8[1] = ms / steps;
    out8[4] = (double)(step_graphs.size() + build_graphs.size());
    out8[5] = (double)n_nodes;
    out8[6] = t_min;
    out8[7] = t_max;
  }
  env->launches += (int64_t)steps * (int64_t)n_nodes;
done:
  if (st) cudaStreamSynchronize(st);
  for (cudaGraphExec_t g : step_graphs) cudaGraphExecDestroy(g);
  for (cudaGraphExec_t g : build_graphs) cudaGraphExecDestroy(g);
  for (int k = 0; k < 4; ++k)
    if (ev[k]) cudaEventDestroy(ev[k]);
  cudaFree(arm0); cudaFree(arm1); cudaFree(base0); cudaFree(base1); cudaFree(valid); cudaFree(step_ctr); cudaFree(acc);
  if (st) cudaStreamDestroy(st);
  return rc;
}
