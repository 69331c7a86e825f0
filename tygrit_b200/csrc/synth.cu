// synth.cu — analytic depth-frame generator for the self-filter workload (benchmark input only).
//
// The reference gets depth images from the ManiSkill/SAPIEN renderer (unavailable: no graphics
// stack).  SURVEY.md §8d replaces it by plain ray casting: each head-camera pixel's ray is
// intersected with the room shell, the furniture cuboids and the robot's own 111 spheres, so that
// robot pixels exist to be masked.  tests/ holds the numpy twin of this arithmetic.
#include <math.h>

#include "fk.cuh"

struct TgSynthCam {
  float r[9], t[3];
  float c[TGM_N_SPHERES * 4];
};

__device__ __forceinline__ unsigned tg_hash(unsigned a, unsigned b, unsigned c) {
  unsigned h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
  return h;
}

__global__ void __launch_bounds__(256)
k_synth_depth(const float* __restrict__ states, const float* __restrict__ boxes, int nb, int H, int W, float fx, float fy,
              float cx, float cy, float half_x, float half_y, float height, unsigned seed, float* __restrict__ depth) {
  __shared__ TgSynthCam cam;
  __shared__ float sbox[64 * 6];
  const int s = blockIdx.y;
  if (threadIdx.x == 0) {
    const float* st = states + 13 * (size_t)s;
    float q[10];
    for (int k = 0; k < 10; ++k) q[k] = st[3 + k];
    TgFrame L[TGM_N_LINKS];
    tg_fk_links(q, L);
    const TgBase b = tg_base(st[0], st[1], st[2]);
    for (int j = 0; j < TGM_N_SPHERES; ++j) {
      const float4 l = TGM_SPH_REF[j];
      const float3 pb = tg_frame_apply(L[TGM_SPH_LINK[j]], l.x, l.y, l.z);
      const float3 pw = tg_to_world(b, pb.x, pb.y, pb.z);
      cam.c[4 * j] = pw.x; cam.c[4 * j + 1] = pw.y; cam.c[4 * j + 2] = pw.z; cam.c[4 * j + 3] = l.w;
    }
    const TgFrame& Tt = L[TGM_HEAD_TILT_LINK];
    const float3 cb = tg_frame_apply(Tt, TGM_HEAD_CAMERA_OFF[0], TGM_HEAD_CAMERA_OFF[1], TGM_HEAD_CAMERA_OFF[2]);
    const float3 cw = tg_to_world(b, cb.x, cb.y, cb.z);
    cam.t[0] = cw.x; cam.t[1] = cw.y; cam.t[2] = cw.z;
    float Rw[9];
    for (int c = 0; c < 3; ++c) {
      Rw[c] = b.c * Tt.r[c] - b.s * Tt.r[3 + c];
      Rw[3 + c] = b.s * Tt.r[c] + b.c * Tt.r[3 + c];
      Rw[6 + c] = Tt.r[6 + c];
    }
    for (int r = 0; r < 3; ++r) {
      cam.r[3 * r + 0] = -Rw[3 * r + 1];
      cam.r[3 * r + 1] = -Rw[3 * r + 2];
      cam.r[3 * r + 2] = Rw[3 * r + 0];
    }
  }
  for (int j = threadIdx.x; j < nb * 6; j += 256) sbox[j] = boxes[j];
  __syncthreads();
  const int pix = blockIdx.x * 256 + threadIdx.x;
  if (pix >= H * W) return;
  const int v = pix / W, u = pix - v * W;
  // ray in camera (OpenCV) axes with unit z: the hit parameter t IS the depth
  const float dxc = ((float)u - cx) / fx, dyc = ((float)v - cy) / fy;
  const float dx = cam.r[0] * dxc + cam.r[1] * dyc + cam.r[2];
  const float dy = cam.r[3] * dxc + cam.r[4] * dyc + cam.r[5];
  const float dz = cam.r[6] * dxc + cam.r[7] * dyc + cam.r[8];
  const float ox = cam.t[0], oy = cam.t[1], oz = cam.t[2];
  float best = 1e30f;
  // room shell (camera is inside): exit distance of the box [-hx,hx] x [-hy,hy] x [0,h]
  {
    const float tx = dx > 0 ? (half_x - ox) / dx : (dx < 0 ? (-half_x - ox) / dx : 1e30f);
    const float ty = dy > 0 ? (half_y - oy) / dy : (dy < 0 ? (-half_y - oy) / dy : 1e30f);
    const float tz = dz > 0 ? (height - oz) / dz : (dz < 0 ? (0.f - oz) / dz : 1e30f);
    const float t = fminf(tx, fminf(ty, tz));
    if (t > 0) best = t;
  }
  for (int b = 0; b < nb; ++b) {  // furniture: slab test
    const float* bx = sbox + 6 * b;
    float t0 = 0.f, t1 = best;
    const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float inv = 1.0f / d[a];
      float ta = (bx[a] - o[a]) * inv, tb = (bx[3 + a] - o[a]) * inv;
      if (ta > tb) { const float tmp = ta; ta = tb; tb = tmp; }
      t0 = fmaxf(t0, ta);
      t1 = fminf(t1, tb);
    }
    if (t0 <= t1 && t0 > 0.f) best = t0;
  }
  const float dd = dx * dx + dy * dy + dz * dz;
  for (int j = 0; j < TGM_N_SPHERES; ++j) {  // the robot's own body
    const float lx = cam.c[4 * j] - ox, ly = cam.c[4 * j + 1] - oy, lz = cam.c[4 * j + 2] - oz, r = cam.c[4 * j + 3];
    const float bq = lx * dx + ly * dy + lz * dz;
    const float cq = lx * lx + ly * ly + lz * lz - r * r;
    const float disc = bq * bq - dd * cq;
    if (disc >= 0.f) {
      const float t = (bq - sqrtf(disc)) / dd;
      if (t > 0.02f && t < best) best = t;
    }
  }
  float out = best < 1e29f ? best : 0.f;
  if ((tg_hash(seed, (unsigned)s, (unsigned)pix) & 1023u) < 20u) out = 0.f;  // ~2 % invalid pixels
  depth[(size_t)s * H * W + pix] = out;
}

extern "C" int tg_synth_depth(const float* states, int64_t S, const float* boxes, int nb, int H, int W, const double* K,
                              float unused, uint64_t seed, float* depth, void* stream) {
  (void)unused;
  TG_CHECK_ARG(S >= 0 && S < 65536 && H > 0 && W > 0 && nb >= 0 && nb <= 64, "tg_synth_depth: bad shape (<= 64 boxes)");
  TG_CHECK_ARG(S == 0 || (states && K && depth && (nb == 0 || boxes)), "tg_synth_depth: NULL buffer");
  if (S == 0) return TG_OK;
  dim3 grid(tg_div_up((int64_t)H * W, 256), (unsigned)S);
  k_synth_depth<<<grid, 256, 0, (cudaStream_t)stream>>>(states, boxes, nb, H, W, (float)K[0], (float)K[4], (float)K[2],
                                                        (float)K[5], 4.0f, 3.0f, 2.6f, (unsigned)seed, depth);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}
