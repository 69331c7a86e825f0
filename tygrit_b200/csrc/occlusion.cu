// occlusion.cu — callers next to the hot path.
//
// * tg_check_occlusion*: batched self-occlusion check.  Reference: TyGrit/checker/occlusion.py:24-116 —
//   the 111 robot spheres and a target sphere are projected into the head camera (pinhole, integer pixel
//   centre and radius), drawn as filled discs with cv2.circle, and the two masks are intersected.  Here
//   one warp handles one robot state: the target disc becomes a per-image-row half-width table in shared
//   memory, every lane walks the scanlines of its robot discs and tests row-interval overlap.  The disc
//   scanlines follow OpenCV's midpoint circle (imgproc/src/drawing.cpp, `Circle`), so the verdict is the
//   reference's pixel-mask verdict, not a geometric approximation.  The arithmetic is fp64 like the
//   reference (truncation to integer pixels makes fp32 flip verdicts); the op is tiny (S x 111 discs).
// * tg_fk_frames_f64: the fp64 FK chain that feeds it (kinematics/fetch/fk_numpy.py:65-155).
// * tg_first_valid: first collision-free candidate per group — the reduction that replaces the serial
//   loop of subgoal_generator/samplers/grasp_sampler.py:109-137 after ONE tg_check_configs call.
#define TGM_WANT_F64
#include "common.cuh"
#include "fetch_model.cuh"

#define TGO_MAX_RADIUS 65536  // discs larger than this (sphere closer than ~0.2 mm to the pinhole) are clamped

struct TgFrameD {
  double r[9];  // row-major rotation
  double t[3];
};

// C = P * Trans(o) * Rot_axis(q);  kind: 0 fixed, 1 prismatic z, 2/3/4 = revolute x/y/z
__device__ void tgo_child(const TgFrameD& P, const double* o, int kind, double q, TgFrameD& C) {
  const double oz = kind == 1 ? o[2] + q : o[2];
  for (int i = 0; i < 3; ++i) C.t[i] = P.r[3 * i] * o[0] + P.r[3 * i + 1] * o[1] + P.r[3 * i + 2] * oz + P.t[i];
  if (kind < 2) {
    for (int i = 0; i < 9; ++i) C.r[i] = P.r[i];
    return;
  }
  double s, c;
  sincos(q, &s, &c);
  const int a = (kind == 2) ? 1 : (kind == 3) ? 2 : 0;
  const int b = (kind == 2) ? 2 : (kind == 3) ? 0 : 1;
  const int k = 3 - a - b;
  for (int i = 0; i < 3; ++i) {
    const double pa = P.r[3 * i + a], pb = P.r[3 * i + b];
    C.r[3 * i + a] = pa * c + pb * s;
    C.r[3 * i + b] = pb * c - pa * s;
    C.r[3 * i + k] = P.r[3 * i + k];
  }
}

// one thread per state: 15 link frames (3x4 row-major, base_link frame) of q10 = torso, 7 arm joints, head pan, head tilt
__global__ void k_fk_frames_f64(const double* __restrict__ q10, int S, double* __restrict__ frames) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  TgFrameD F[TGM_N_LINKS];
  for (int i = 0; i < 9; ++i) F[0].r[i] = (i % 4 == 0) ? 1.0 : 0.0;
  F[0].t[0] = F[0].t[1] = F[0].t[2] = 0.0;
  for (int i = 1; i < TGM_N_LINKS; ++i) {
    const int slot = TGM_LINK_QSLOT[i];
    tgo_child(F[TGM_LINK_PARENT[i]], TGM_LINK_OFF_D[i], TGM_LINK_KIND[i], slot >= 0 ? q10[10 * (size_t)s + slot] : 0.0, F[i]);
  }
  double* out = frames + (size_t)s * TGM_N_LINKS * 12;
  for (int i = 0; i < TGM_N_LINKS; ++i)
    for (int r = 0; r < 3; ++r) {
      out[12 * i + 4 * r + 0] = F[i].r[3 * r + 0];
      out[12 * i + 4 * r + 1] = F[i].r[3 * r + 1];
      out[12 * i + 4 * r + 2] = F[i].r[3 * r + 2];
      out[12 * i + 4 * r + 3] = F[i].t[r];
    }
}

extern "C" int tg_fk_frames_f64(const double* q10, int64_t S, double* frames, void* stream) {
  TG_CHECK_ARG(S >= 0 && S < (1ll << 31), "tg_fk_frames_f64: S out of range");
  TG_CHECK_ARG(S == 0 || (q10 && frames), "tg_fk_frames_f64: NULL buffer");
  if (S == 0) return TG_OK;
  k_fk_frames_f64<<<tg_div_up(S, 64), 64, 0, (cudaStream_t)stream>>>(q10, (int)S, frames);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}

// Python's int(): truncation toward zero (clamped so that the cast is defined)
__device__ __forceinline__ long long tgo_int(double x) {
  if (!(x == x)) return 0x7fffffffffffll;  // NaN never lands in the image
  return (long long)fmin(fmax(x, -4.0e15), 4.0e15);
}

struct TgDisc {
  int u, v, r;
  bool drawn;
};

// occlusion.py:38-43: skipped when behind the camera or when the truncated centre is outside the image
__device__ __forceinline__ TgDisc tgo_project(double X, double Y, double Z, double radius, double fx, double fy, double cx,
                                              double cy, int H, int W) {
  TgDisc d;
  d.u = d.v = d.r = 0;
  d.drawn = false;
  if (!(Z > 0.0)) return d;
  const long long u = tgo_int(fx * X / Z + cx), v = tgo_int(fy * Y / Z + cy);
  long long r = tgo_int(fx * radius / Z);
  if (r < 1) r = 1;
  if (r > TGO_MAX_RADIUS) r = TGO_MAX_RADIUS;
  if (u < 0 || u >= W || v < 0 || v >= H) return d;
  d.u = (int)u; d.v = (int)v; d.r = (int)r;
  d.drawn = true;
  return d;
}

// OpenCV's filled circle: every step of the midpoint walk fills rows cy -+ dy with half-width dx and rows
// cy -+ dx with half-width dy.  `emit(row_offset, half_width)` returns true to stop the walk.
template <class F>
__device__ __forceinline__ void tgo_walk_disc(int radius, F emit) {
  int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
  while (dx >= dy) {
    if (emit(dy, dx) || emit(dx, dy)) return;
    ++dy;
    err += plus;
    plus += 2;
    if (err > 0) {
      err -= minus;
      --dx;
      minus -= 2;
    }
  }
}

// one warp per state.  frames: S x 15 x 12 doubles; target: xyzr in base_link (stride 0 = one target for all states)
__global__ void __launch_bounds__(128)
k_occlusion(const double* __restrict__ frames, uint32_t present, const double* __restrict__ target, int target_stride, int S,
            double fx, double fy, double cx, double cy, int H, int W, uint8_t* __restrict__ out) {
  extern __shared__ int tgo_rows[];  // [4 warps][H]: half-width of the target disc in image row y, -1 = none
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int s = blockIdx.x * 4 + w;
  if (s >= S) return;  // warp-uniform
  int* hw = tgo_rows + (size_t)w * H;
  if (!(present >> TGM_HEAD_TILT_LINK & 1u)) {  // occlusion.py:70-72
    if (lane == 0) out[s] = 0;
    return;
  }
  const double* Fs = frames + (size_t)s * TGM_N_LINKS * 12;
  // camera_link pose in base_link: rotation of head_tilt, origin moved by the camera offset (occlusion.py:75-76)
  const double* Th = Fs + 12 * TGM_HEAD_TILT_LINK;
  double R[9], tc[3];
  for (int r = 0; r < 3; ++r) {
    R[3 * r] = Th[4 * r]; R[3 * r + 1] = Th[4 * r + 1]; R[3 * r + 2] = Th[4 * r + 2];
    tc[r] = Th[4 * r + 3] + (Th[4 * r] * TGM_HEAD_CAMERA_OFF_D[0] + Th[4 * r + 1] * TGM_HEAD_CAMERA_OFF_D[1] +
                             Th[4 * r + 2] * TGM_HEAD_CAMERA_OFF_D[2]);
  }
  // base_link point -> OpenCV camera: p_link = R^T (p - tc); cv = (-p_link.y, -p_link.z, p_link.x)  (occlusion.py:79-83)
  auto to_cv = [&](double px, double py, double pz, double& X, double& Y, double& Z) {
    const double dx = px - tc[0], dy = py - tc[1], dz = pz - tc[2];
    const double lx = R[0] * dx + R[3] * dy + R[6] * dz;
    const double ly = R[1] * dx + R[4] * dy + R[7] * dz;
    const double lz = R[2] * dx + R[5] * dy + R[8] * dz;
    X = -ly; Y = -lz; Z = lx;
  };
  const double* tg = target + (size_t)s * target_stride;
  double X, Y, Z;
  to_cv(tg[0], tg[1], tg[2], X, Y, Z);
  const TgDisc T = tgo_project(X, Y, Z, tg[3], fx, fy, cx, cy, H, W);
  if (!T.drawn) {  // empty target mask
    if (lane == 0) out[s] = 0;
    return;
  }
  for (int y = lane; y < H; y += 32) hw[y] = -1;
  __syncwarp();
  if (lane == 0)
    tgo_walk_disc(T.r, [&](int off, int half) {
      const int y0 = T.v - off, y1 = T.v + off;
      if (y0 >= 0 && y0 < H) hw[y0] = max(hw[y0], half);
      if (y1 >= 0 && y1 < H) hw[y1] = max(hw[y1], half);
      return false;
    });
  __syncwarp();
  bool hit = false;
  for (int j = lane; j < TGM_N_SPHERES && !hit; j += 32) {
    const int link = TGM_SPH_LINK[j];
    if (!(present >> link & 1u)) continue;  // occlusion.py:91-93
    const double* F = Fs + 12 * link;
    const double* sp = TGM_SPH_REF_D[j];
    const double bx = F[0] * sp[0] + F[1] * sp[1] + F[2] * sp[2] + F[3];
    const double by = F[4] * sp[0] + F[5] * sp[1] + F[6] * sp[2] + F[7];
    const double bz = F[8] * sp[0] + F[9] * sp[1] + F[10] * sp[2] + F[11];
    to_cv(bx, by, bz, X, Y, Z);
    const TgDisc D = tgo_project(X, Y, Z, sp[3], fx, fy, cx, cy, H, W);
    if (!D.drawn) continue;
    if (abs(D.v - T.v) > D.r + T.r || abs(D.u - T.u) > D.r + T.r) continue;
    tgo_walk_disc(D.r, [&](int off, int half) {
      const int a0 = max(D.u - half, 0), a1 = min(D.u + half, W - 1);
      for (int sgn = 0; sgn < 2; ++sgn) {
        const int y = sgn ? D.v + off : D.v - off;
        if (y < 0 || y >= H) continue;
        const int h = hw[y];
        if (h < 0) continue;
        if (max(a0, max(T.u - h, 0)) <= min(a1, min(T.u + h, W - 1))) hit = true;
      }
      return hit;
    });
  }
  hit = __any_sync(0xffffffffu, hit);
  if (lane == 0) out[s] = hit ? 1 : 0;
}

extern "C" int tg_check_occlusion_frames(const double* frames, uint32_t links_present, const double* target, int64_t n_targets,
                                         int64_t S, const double* K, int H, int W, uint8_t* occluded, void* stream) {
  TG_CHECK_ARG(S >= 0 && S < (1ll << 29), "tg_check_occlusion: S out of range");
  TG_CHECK_ARG(H > 0 && W > 0 && H <= 8192 && W <= 65536, "tg_check_occlusion: image shape out of range (H <= 8192)");
  TG_CHECK_ARG(n_targets == 1 || n_targets == S, "tg_check_occlusion: one target, or one per state");
  TG_CHECK_ARG(S == 0 || (frames && target && K && occluded), "tg_check_occlusion: NULL buffer");
  if (S == 0) return TG_OK;
  const size_t smem = sizeof(int) * 4 * (size_t)H;
  if (smem > 48 * 1024) TG_CUDA(cudaFuncSetAttribute(k_occlusion, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_occlusion<<<tg_div_up(S, 4), 128, smem, (cudaStream_t)stream>>>(frames, links_present, target, n_targets == 1 ? 0 : 4, (int)S,
                                                                   K[0], K[4], K[2], K[5], H, W, occluded);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}

extern "C" int tg_check_occlusion(const double* q10, const double* target, int64_t n_targets, int64_t S, const double* K, int H,
                                  int W, uint8_t* occluded, void* stream) {
  TG_CHECK_ARG(S >= 0 && S < (1ll << 29), "tg_check_occlusion: S out of range");
  TG_CHECK_ARG(S == 0 || q10, "tg_check_occlusion: NULL buffer");
  if (S == 0) return TG_OK;
  double* frames = nullptr;
  TG_CUDA(cudaMallocAsync(&frames, sizeof(double) * 12 * TGM_N_LINKS * (size_t)S, (cudaStream_t)stream));
  int rc = tg_fk_frames_f64(q10, S, frames, stream);
  if (rc == TG_OK) rc = tg_check_occlusion_frames(frames, (1u << TGM_N_LINKS) - 1u, target, n_targets, S, K, H, W, occluded, stream);
  cudaFreeAsync(frames, (cudaStream_t)stream);
  return rc;
}

// ---- first valid candidate per group ----------------------------------------------------------------------------------
__global__ void k_first_valid(const uint8_t* __restrict__ valid, long long B, int group, int* __restrict__ first) {
  const long long g = blockIdx.x;  // one block per group
  const uint8_t* v = valid + g * group;
  const int n = (int)min((long long)group, B - g * group);
  __shared__ int best;
  if (threadIdx.x == 0) best = 0x7fffffff;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += blockDim.x) {  // stop after the first stripe that holds a valid candidate
    const int i = i0 + threadIdx.x;
    if (i < n && v[i]) atomicMin(&best, i);
    __syncthreads();
    if (best != 0x7fffffff) break;
  }
  if (threadIdx.x == 0) first[g] = best == 0x7fffffff ? -1 : best;
}

extern "C" int tg_first_valid(const uint8_t* valid, int64_t B, int group, int32_t* first, void* stream) {
  TG_CHECK_ARG(B >= 0 && group >= 1, "tg_first_valid: bad sizes");
  TG_CHECK_ARG(B == 0 || (valid && first), "tg_first_valid: NULL buffer");
  if (B == 0) return TG_OK;
  const long long groups = (B + group - 1) / group;
  TG_CHECK_ARG(groups < (1ll << 31), "tg_first_valid: too many groups");
  k_first_valid<<<(unsigned)groups, 128, 0, (cudaStream_t)stream>>>(valid, B, group, first);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}
