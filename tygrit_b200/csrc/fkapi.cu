// fkapi.cu — batched FK entry points (parity / BatchFKSolver surface).
//
// tg_fk_frames replaces TyGrit/kinematics/fetch/fk_torch.py:119-283 (batch_forward_kinematics,
// ~25 bmm + dozens of small allocations per call) with one kernel: one lane per configuration,
// the 15 link frames stay in registers and are written once.
// tg_fk_spheres produces the 111 world sphere centres, which the reference only ever computes
// per link for a single configuration (kinematics/robot_filter.py:35-44).
#include "fk.cuh"

__global__ void __launch_bounds__(128) k_fk_frames(const float* __restrict__ q10, int B, float* __restrict__ frames) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  float q[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) q[k] = q10[10 * (size_t)i + k];
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q, L);
  float4* out = reinterpret_cast<float4*>(frames + (size_t)i * TGM_N_LINKS * 16);
#pragma unroll
  for (int l = 0; l < TGM_N_LINKS; ++l) {
    out[4 * l + 0] = make_float4(L[l].r[0], L[l].r[1], L[l].r[2], L[l].t[0]);
    out[4 * l + 1] = make_float4(L[l].r[3], L[l].r[4], L[l].r[5], L[l].t[1]);
    out[4 * l + 2] = make_float4(L[l].r[6], L[l].r[7], L[l].r[8], L[l].t[2]);
    out[4 * l + 3] = make_float4(0.f, 0.f, 0.f, 1.f);
  }
}

__global__ void __launch_bounds__(128)
k_fk_spheres(const float* __restrict__ arm, const float* __restrict__ base, const float* __restrict__ head2, int B,
             float* __restrict__ centres) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  float q[10];
#pragma unroll
  for (int k = 0; k < 8; ++k) q[k] = arm[8 * (size_t)i + k];
  q[8] = head2 ? head2[2 * (size_t)i] : 0.f;
  q[9] = head2 ? head2[2 * (size_t)i + 1] : 0.f;
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q, L);
  const TgBase b = tg_base(base[3 * (size_t)i], base[3 * (size_t)i + 1], base[3 * (size_t)i + 2]);
  float* out = centres + (size_t)i * TGM_N_SPHERES * 3;
  for (int j = 0; j < TGM_N_SPHERES; ++j) {
    const float4 l = TGM_SPH_REF[j];
    const float3 pb = tg_frame_apply(L[TGM_SPH_LINK[j]], l.x, l.y, l.z);
    const float3 pw = tg_to_world(b, pb.x, pb.y, pb.z);
    out[3 * j] = pw.x;
    out[3 * j + 1] = pw.y;
    out[3 * j + 2] = pw.z;
  }
}

extern "C" int tg_fk_frames(const float* q10, int64_t B, float* frames, void* stream) {
  TG_CHECK_ARG(B >= 0 && (B == 0 || (q10 && frames)), "tg_fk_frames: NULL buffer");
  TG_CHECK_ARG(B < (1ll << 31), "tg_fk_frames: batch too large");
  TG_CHECK_ARG(((uintptr_t)frames & 15) == 0, "tg_fk_frames: frames must be 16-byte aligned");
  if (B == 0) return TG_OK;
  k_fk_frames<<<tg_div_up(B, 128), 128, 0, (cudaStream_t)stream>>>(q10, (int)B, frames);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}

extern "C" int tg_fk_spheres(const float* arm, const float* base, const float* head2, int64_t B, float* centres, void* stream) {
  TG_CHECK_ARG(B >= 0 && (B == 0 || (arm && base && centres)), "tg_fk_spheres: NULL buffer");
  TG_CHECK_ARG(B < (1ll << 31), "tg_fk_spheres: batch too large");
  if (B == 0) return TG_OK;
  k_fk_spheres<<<tg_div_up(B, 128), 128, 0, (cudaStream_t)stream>>>(arm, base, head2, (int)B, centres);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}
