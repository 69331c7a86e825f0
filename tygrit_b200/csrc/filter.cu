// filter.cu — K4: robot self-filter (points inside the robot's collision volume are masked out).
//
// Reference: TyGrit/kinematics/robot_filter.py:13-64 (uniform radius for all 111 spheres, strict
// `d2 < r2` removes, centres cast to f32 before the test) and the depth back-projection of
// TyGrit/utils/depth.py:13-84.  HBM-bound: 4 B depth in + 1 B mask out per (state, pixel); the 111
// centres of a state live in shared memory and a bounding box rejects non-robot pixels early.
//
// The inner test is written with explicit round-to-nearest mul/add (no FMA contraction) in the
// order ((dx*dx + dy*dy) + dz*dz), which is what numpy evaluates; given identical f32 centres the
// mask is bit-identical to the reference, not just equal outside a tolerance band.
#include <math.h>

#include "fk.cuh"

#define TGF_BLOCK 256
#define TGF_MAX_CENTRES 1024

struct TgFilterState {  // per robot state, produced by k_filter_prepare
  float cam_r[9];       // cam2world rotation (OpenCV camera axes -> world)
  float cam_t[3];
  float lo[3], hi[3];   // bounding box of the sphere centres
  float c[TGM_N_SPHERES * 3];
  float cc[TGM_N_SPHERES * 3];  // the same centres in the camera frame (cull tests only, never the verdict)
  float img_box[4];             // conservative pixel box (u0, v0, u1, v1) of every sphere that can be seen at all
};

__device__ __forceinline__ bool tg_inside_any(const float* __restrict__ c, int S, float px, float py, float pz, float r2) {
  bool inside = false;
  for (int s = 0; s < S; ++s) {
    const float dx = __fsub_rn(px, c[3 * s]), dy = __fsub_rn(py, c[3 * s + 1]), dz = __fsub_rn(pz, c[3 * s + 2]);
    const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    inside |= d2 < r2;
  }
  return inside;
}

// ---- generic points x centres ---------------------------------------------------------------
__global__ void __launch_bounds__(TGF_BLOCK)
k_filter_points(const float* __restrict__ pts, int N, const float* __restrict__ centres, int S, float r2, float r,
                uint8_t* __restrict__ keep) {
  __shared__ float sc[TGF_MAX_CENTRES * 3];
  __shared__ float sbox[6];
  for (int j = threadIdx.x; j < 3 * S; j += TGF_BLOCK) sc[j] = centres[j];
  __syncthreads();
  if (threadIdx.x < 3) {
    float lo = INFINITY, hi = -INFINITY;
    for (int s = 0; s < S; ++s) {
      lo = fminf(lo, sc[3 * s + threadIdx.x]);
      hi = fmaxf(hi, sc[3 * s + threadIdx.x]);
    }
    sbox[threadIdx.x] = lo - r - 1e-3f;
    sbox[3 + threadIdx.x] = hi + r + 1e-3f;
  }
  __syncthreads();
  for (int i = blockIdx.x * TGF_BLOCK + threadIdx.x; i < N; i += gridDim.x * TGF_BLOCK) {
    const float px = pts[3 * (size_t)i], py = pts[3 * (size_t)i + 1], pz = pts[3 * (size_t)i + 2];
    bool inside = false;
    if (px >= sbox[0] && px <= sbox[3] && py >= sbox[1] && py <= sbox[4] && pz >= sbox[2] && pz <= sbox[5])
      inside = tg_inside_any(sc, S, px, py, pz, r2);
    keep[i] = inside ? 0 : 1;
  }
}

extern "C" int tg_filter_points(const float* points, int64_t N, const float* centres, int S, float radius, uint8_t* keep,
                                void* stream) {
  TG_CHECK_ARG(N >= 0 && N < (1ll << 31), "tg_filter_points: N out of range");
  TG_CHECK_ARG(S >= 0 && S <= TGF_MAX_CENTRES, "tg_filter_points: at most 1024 centres");
  TG_CHECK_ARG(N == 0 || (points && keep && (S == 0 || centres)), "tg_filter_points: NULL buffer");
  if (N == 0) return TG_OK;
  const float r2 = (float)((double)radius * (double)radius);  // python: r2 = r * r (f64), compared in f32
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int blocks = min(tg_div_up(N, TGF_BLOCK), sms * 8);
  k_filter_points<<<blocks, TGF_BLOCK, 0, (cudaStream_t)stream>>>(points, (int)N, centres, S, r2, radius, keep);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}

// ---- batched depth self-filter --------------------------------------------------------------
struct TgDepthParams {
  int H, W, stride, Hs, Ws;
  double fx, fy, cx, cy;  // intrinsics of the decimated image (K rows 0,1 divided by stride in f64)
  float zmin, zmax, r2;
};

// One warp per robot state: lane 0 runs the FK chain and the camera pose, then the lanes share the
// 111 spheres (world centres, camera-frame centres, bounding boxes).
__global__ void __launch_bounds__(128)
k_filter_prepare(const float* __restrict__ states, int S, float radius, TgDepthParams p, TgFilterState* __restrict__ out) {
  __shared__ TgFrame sL[4][TGM_N_LINKS];
  __shared__ TgBase sB[4];
  __shared__ float sCam[4][12];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int s = blockIdx.x * 4 + w;
  if (s >= S) return;  // warp-uniform
  TgFilterState& o = out[s];
  if (lane == 0) {
    const float* st = states + 13 * (size_t)s;
    float q[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) q[k] = st[3 + k];
    tg_fk_links(q, sL[w]);
    const TgBase b = tg_base(st[0], st[1], st[2]);
    sB[w] = b;
    // cam2world = SE2(base) * T(head_tilt) * Trans(cam offset) * R_cv->link   (camera.py:25-58)
    const TgFrame& Tt = sL[w][TGM_HEAD_TILT_LINK];
    const float3 cb = tg_frame_apply(Tt, TGM_HEAD_CAMERA_OFF[0], TGM_HEAD_CAMERA_OFF[1], TGM_HEAD_CAMERA_OFF[2]);
    const float3 cw = tg_to_world(b, cb.x, cb.y, cb.z);
    sCam[w][9] = cw.x; sCam[w][10] = cw.y; sCam[w][11] = cw.z;
    // world rotation of head_tilt: Rz(theta) * R ; then columns: cv x -> -link y, cv y -> -link z, cv z -> link x
    float Rw[9];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      Rw[c] = fmaf(b.c, Tt.r[c], -b.s * Tt.r[3 + c]);
      Rw[3 + c] = fmaf(b.s, Tt.r[c], b.c * Tt.r[3 + c]);
      Rw[6 + c] = Tt.r[6 + c];
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      sCam[w][3 * r + 0] = -Rw[3 * r + 1];
      sCam[w][3 * r + 1] = -Rw[3 * r + 2];
      sCam[w][3 * r + 2] = Rw[3 * r + 0];
    }
  }
  __syncwarp();
  if (lane < 9) o.cam_r[lane] = sCam[w][lane];
  else if (lane < 12) o.cam_t[lane - 9] = sCam[w][lane];
  const float* R = sCam[w];
  const float* t = sCam[w] + 9;
  const TgBase b = sB[w];
  const float rp = radius + 2e-3f;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float ib[4] = {INFINITY, INFINITY, -INFINITY, -INFINITY};
  for (int j = lane; j < TGM_N_SPHERES; j += 32) {
    const float4 l = TGM_SPH_REF[j];
    const float3 pb = tg_frame_apply(sL[w][TGM_SPH_LINK[j]], l.x, l.y, l.z);
    const float3 pw = tg_to_world(b, pb.x, pb.y, pb.z);
    o.c[3 * j] = pw.x; o.c[3 * j + 1] = pw.y; o.c[3 * j + 2] = pw.z;
    lo[0] = fminf(lo[0], pw.x); hi[0] = fmaxf(hi[0], pw.x);
    lo[1] = fminf(lo[1], pw.y); hi[1] = fmaxf(hi[1], pw.y);
    lo[2] = fminf(lo[2], pw.z); hi[2] = fmaxf(hi[2], pw.z);
    // camera frame: c_cam = R^T (c_world - t)
    const float dx = pw.x - t[0], dy = pw.y - t[1], dz = pw.z - t[2];
    float cc[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) cc[a] = fmaf(R[a], dx, fmaf(R[3 + a], dy, R[6 + a] * dz));
    o.cc[3 * j] = cc[0]; o.cc[3 * j + 1] = cc[1]; o.cc[3 * j + 2] = cc[2];
    // pixel box of everything this sphere can contain: depth in [zl, zh], |x - X| < rp, |y - Y| < rp
    const float zl = fmaxf(cc[2] - rp, p.zmin - 1e-3f), zh = fminf(cc[2] + rp, p.zmax + 1e-3f);
    if (zl <= zh && zh > 0.f) {
      if (zl < 0.01f) {  // reaches the camera: no useful bound
        ib[0] = ib[1] = -INFINITY;
        ib[2] = ib[3] = INFINITY;
      } else {
        const float fx = (float)p.fx, fy = (float)p.fy, cx = (float)p.cx, cy = (float)p.cy;
        const float xa = cc[0] - rp, xb = cc[0] + rp, ya = cc[1] - rp, yb = cc[1] + rp;
        ib[0] = fminf(ib[0], cx + fx * (xa >= 0.f ? xa / zh : xa / zl) - 1.5f);
        ib[2] = fmaxf(ib[2], cx + fx * (xb >= 0.f ? xb / zl : xb / zh) + 1.5f);
        ib[1] = fminf(ib[1], cy + fy * (ya >= 0.f ? ya / zh : ya / zl) - 1.5f);
        ib[3] = fmaxf(ib[3], cy + fy * (yb >= 0.f ? yb / zl : yb / zh) + 1.5f);
      }
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a)
    for (int of = 16; of; of >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], of));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], of));
    }
  for (int of = 16; of; of >>= 1) {
    ib[0] = fminf(ib[0], __shfl_xor_sync(0xffffffffu, ib[0], of));
    ib[1] = fminf(ib[1], __shfl_xor_sync(0xffffffffu, ib[1], of));
    ib[2] = fmaxf(ib[2], __shfl_xor_sync(0xffffffffu, ib[2], of));
    ib[3] = fmaxf(ib[3], __shfl_xor_sync(0xffffffffu, ib[3], of));
  }
  if (lane < 3) {
    o.lo[lane] = lo[lane] - radius - 1e-3f;
    o.hi[lane] = hi[lane] + radius + 1e-3f;
  }
  if (lane < 4) o.img_box[lane] = ib[lane];
}

// one thread = PIX consecutive decimated pixels of one state
template <int PIX>
__global__ void __launch_bounds__(TGF_BLOCK)
k_self_filter_depth(const float* __restrict__ depth, const TgFilterState* __restrict__ prep, TgDepthParams p, float radius,
                    uint8_t* __restrict__ mask, float* __restrict__ world) {
  __shared__ TgFilterState st;
  const float rp = radius + 2e-3f, rout2 = rp * rp, rin = fmaxf(radius - 2e-3f, 0.f), rin2 = rin * rin;
  const int s = blockIdx.y;
  {
    const float* src = reinterpret_cast<const float*>(prep + s);
    float* dst = reinterpret_cast<float*>(&st);
    for (int j = threadIdx.x; j < (int)(sizeof(TgFilterState) / 4); j += TGF_BLOCK) dst[j] = src[j];
  }
  __syncthreads();
  const int npx = p.Hs * p.Ws;
  // mask-only calls: the spheres that can meet this block's image rows at all (depth interval x projected y-range,
  // conservative), compacted into shared memory — the per-pixel loop below walks ~20 of them instead of 111
  __shared__ float4 s_list[TGM_N_SPHERES];
  __shared__ int s_cnt;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  if (world == nullptr && threadIdx.x < TGM_N_SPHERES) {
    const bool whole = (long long)gridDim.x * TGF_BLOCK * PIX < npx;  // the block loops over several row ranges
    const int p0 = blockIdx.x * TGF_BLOCK * PIX, p1 = min(p0 + TGF_BLOCK * PIX, npx) - 1;
    const int v0 = whole ? 0 : p0 / p.Ws, v1 = whole ? p.Hs - 1 : p1 / p.Ws;
    const float b0 = ((float)v0 - 0.5f - (float)p.cy) * (float)(1.0 / p.fy), b1 = ((float)v1 + 0.5f - (float)p.cy) * (float)(1.0 / p.fy);
    const int j = threadIdx.x;
    const float X = st.cc[3 * j], Y = st.cc[3 * j + 1], Z = st.cc[3 * j + 2];
    const float zl = fmaxf(fmaxf(Z - rp, p.zmin - 1e-3f), 0.f), zh = fminf(Z + rp, p.zmax + 1e-3f);
    if (zl <= zh) {
      const float ymin = (b0 >= 0.f ? zl : zh) * b0, ymax = (b1 >= 0.f ? zh : zl) * b1;
      if (ymin <= Y + rp && ymax >= Y - rp) s_list[atomicAdd(&s_cnt, 1)] = make_float4(X, Y, Z, 0.f);
    }
  }
  __syncthreads();
  const int cnt = s_cnt;
  // a block walks the frame in steps of gridDim.x * TGF_BLOCK * PIX pixels (one step with the launch shape used today)
  for (int first = (blockIdx.x * TGF_BLOCK + threadIdx.x) * PIX; first < npx; first += gridDim.x * TGF_BLOCK * PIX) {
  const float* img = depth + (size_t)s * p.H * p.W;
  float d[PIX];
  if (PIX == 4 && p.stride == 1 && (p.W & 3) == 0) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(img + first));
    d[0] = v.x; d[1 % PIX] = v.y; d[2 % PIX] = v.z; d[3 % PIX] = v.w;
  } else {
#pragma unroll
    for (int k = 0; k < PIX; ++k) {
      const int pix = first + k;
      if (pix < npx) {
        const int v = pix / p.Ws, u = pix - v * p.Ws;
        d[k] = __ldg(img + (size_t)(v * p.stride) * p.W + u * p.stride);
      } else d[k] = 0.f;
    }
  }
  uint8_t out[PIX];
#pragma unroll
  for (int k = 0; k < PIX; ++k) {
    const int pix = first + k;
    const int v = pix / p.Ws, u = pix - v * p.Ws;
    float z = d[k];
    if (isnan(z)) z = 0.f;                       // np.nan_to_num
    else if (isinf(z)) z = z > 0 ? 3.4028234663852886e38f : -3.4028234663852886e38f;
    bool ok = pix < npx && z > 0.f && z >= p.zmin && z <= p.zmax;
    float wx = 0.f, wy = 0.f, wz = 0.f;
    if (ok && world == nullptr) {
      // mask only: settle the pixel in the camera frame where that is safe (see k_self_filter_stream): outside the
      // robot's image-space box, or farther than r + 2 mm from every sphere -> kept; closer than r - 2 mm to one ->
      // removed; only the 4 mm shells take the reference-exact path below
      if ((float)u < st.img_box[0] || (float)u > st.img_box[2] || (float)v < st.img_box[1] || (float)v > st.img_box[3]) {
        out[k] = 1;
        continue;
      }
      const float xa = ((float)u - (float)p.cx) * (float)(1.0 / p.fx) * z, ya = ((float)v - (float)p.cy) * (float)(1.0 / p.fy) * z;
      bool in = false, shell = false;
      for (int s2 = 0; s2 < cnt; ++s2) {
        const float4 e = s_list[s2];
        const float dz = z - e.z;
        if (fabsf(dz) < rp) {
          const float dx = xa - e.x, dy = ya - e.y;
          const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
          in |= d2 < rin2;
          shell |= d2 < rout2;
        }
      }
      if (in || !shell) {
        out[k] = in ? 0 : 1;
        continue;
      }
    }
    if (ok) {
      // depth.py:41-46 evaluated as numpy 2 does: (f32 index - f64 c) * f32 depth / f64 f, rounded once
      const float xc = (float)((((double)(float)u - p.cx) * (double)z) / p.fx);
      const float yc = (float)((((double)(float)v - p.cy) * (double)z) / p.fy);
      // depth.py:79-83: p_w = p_c @ R^T + t in f32
      wx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[0]), __fmul_rn(yc, st.cam_r[1])), __fmul_rn(z, st.cam_r[2])), st.cam_t[0]);
      wy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[3]), __fmul_rn(yc, st.cam_r[4])), __fmul_rn(z, st.cam_r[5])), st.cam_t[1]);
      wz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[6]), __fmul_rn(yc, st.cam_r[7])), __fmul_rn(z, st.cam_r[8])), st.cam_t[2]);
      if (wx >= st.lo[0] && wx <= st.hi[0] && wy >= st.lo[1] && wy <= st.hi[1] && wz >= st.lo[2] && wz <= st.hi[2])
        ok = !tg_inside_any(st.c, TGM_N_SPHERES, wx, wy, wz, p.r2);
    }
    out[k] = ok ? 1 : 0;
    if (world != nullptr && pix < npx) {
      float* w = world + ((size_t)s * npx + pix) * 3;
      w[0] = wx; w[1] = wy; w[2] = wz;
    }
  }
  uint8_t* m = mask + (size_t)s * npx + first;
  if (PIX == 4 && first + 3 < npx && ((((size_t)s * npx + first) & 3) == 0)) {
    *reinterpret_cast<uint32_t*>(m) = (uint32_t)out[0] | ((uint32_t)out[1 % PIX] << 8) | ((uint32_t)out[2 % PIX] << 16) | ((uint32_t)out[3 % PIX] << 24);
  } else {
#pragma unroll
    for (int k = 0; k < PIX; ++k)
      if (first + k < npx) m[k] = out[k];
  }
  }
}

// ---- K4 v3: streaming fast path (full-resolution frames, mask only) --------------------------------
// Persistent blocks, four per SM, each one producer warp + four consumer warps.  A block owns every gridDim-th
// "chunk" (R = 8 full image rows of one state = one contiguous 20 KB piece of the depth array) and pulls them
// through a 2-slot shared-memory ring with 1-D bulk async copies (cp.async.bulk + mbarrier complete_tx, L2
// evict-first): 8 x 20 KB per SM in flight, independent of register pressure and occupancy, which is what an
// HBM-bound stream needs (6.5 TB/s x ~1 us latency = 44 KB per SM at the very least).  While a chunk is in flight
// the producer warp culls the robot's spheres against the chunk's rows (band list); consumers read 2 x 4 pixels
// per lane from the ring (conflict-free LDS.128); a pixel costs one integer compare unless its 32 x 8 tile has
// candidates in the band list AND in its depth range — then it is classified in the camera frame, and only pixels
// in a 4 mm shell around a sphere surface are transformed and tested with the generic kernel's arithmetic (same
// rounding, same verdict), so the mask stays byte-identical.  Measured A/B (C3, ms): 2 blocks x 8 consumers, 32 x 4
// tiles, 4 slots 0.389; 32 x 8 tiles with 10 consumers 0.42, 16-row chunks 0.35, 3 x 5 consumers 0.346, 4 x 4
// consumers with 2 slots 0.334 (more producers, more blocks to hide the fixed-latency dependencies).
#ifndef TGS_STAGES
#define TGS_STAGES 2
#endif
#ifndef TGS_STAGE_BYTES
#define TGS_STAGE_BYTES 20480
#endif
#ifndef TGS_MAX_ROWS
#define TGS_MAX_ROWS 8
#endif

__device__ __forceinline__ uint32_t tg_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tg_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tg_smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tg_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tg_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tg_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                   tg_smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(tg_smem_addr(bar)), "l"(policy)
               : "memory");
}
__device__ __forceinline__ bool tg_mbar_try(uint32_t bar_addr, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
               : "=r"(done)
               : "r"(bar_addr), "r"(parity)
               : "memory");
  return done != 0;
}
// spinning warps take issue slots from the warps that have work: back off between polls
__device__ __forceinline__ void tg_mbar_wait(uint32_t bar_addr, uint32_t parity, unsigned backoff_ns) {
  while (!tg_mbar_try(bar_addr, parity)) __nanosleep(backoff_ns);
}

#ifndef TGS_CONSUMERS
#define TGS_CONSUMERS 4                         // consumer warps; warp TGS_CONSUMERS is the producer
#endif
#ifndef TGS_BLOCKS_PER_SM
#define TGS_BLOCKS_PER_SM 4
#endif
#ifndef TGS_NF4
#define TGS_NF4 2                               // float4 per lane and tile: tiles are 32 x 4 or 32 x 8 pixels
#endif
#define TGS_THREADS ((TGS_CONSUMERS + 1) * 32)
#define TGS_STATE_BYTES ((int)((sizeof(TgFilterState) + 127) / 128 * 128))
#define TGS_LIST_BYTES (112 * 16)               // band list: (X, Y, Z, sphere index) of the spheres that can meet the chunk's rows
#define TGS_META_BYTES 128                      // {count, first tile column, last tile column}
// one ring slot = the chunk's depth rows + its robot state + the chunk's band list
#define TGS_SLOT_BYTES (TGS_STAGE_BYTES + TGS_STATE_BYTES + TGS_LIST_BYTES + TGS_META_BYTES)
static_assert(sizeof(TgFilterState) % 16 == 0, "bulk copies move multiples of 16 bytes");

__global__ void __launch_bounds__(TGS_THREADS, TGS_BLOCKS_PER_SM)
k_self_filter_stream(const float* __restrict__ depth, const TgFilterState* __restrict__ prep, TgDepthParams p, float radius,
                     uint8_t* __restrict__ mask, int R, int cpf, int total_chunks) {
  extern __shared__ __align__(128) unsigned char tgs_smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(tgs_smem + (size_t)TGS_STAGES * TGS_SLOT_BYTES);  // full[], empty[]
  const uint32_t full0 = tg_smem_addr(bars), empty0 = full0 + 8 * TGS_STAGES;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // chunk i of this block is chunk blockIdx.x + i * gridDim.x: at any time the grid reads one contiguous window of
  // the depth array, and the frames in which the robot fills the view (far more work) are spread over all blocks
  const int G = (int)gridDim.x;
  const int n = (int)blockIdx.x < total_chunks ? (total_chunks - (int)blockIdx.x + G - 1) / G : 0;
  const int dS = G / cpf, dK = G - dS * cpf;
  int s = (int)blockIdx.x / cpf, kc = (int)blockIdx.x - s * cpf;  // state and chunk-in-frame of chunk i, kept incrementally
  const int W = p.W, H = p.H, w4 = W >> 2, ntx = (W + 31) >> 5;
  const float rp = radius + 2e-3f;
  const float ifx = (float)(1.0 / p.fx), ify = (float)(1.0 / p.fy), cxf = (float)p.cx, cyf = (float)p.cy;
  if (tid == 0) {
#pragma unroll
    for (int k = 0; k < TGS_STAGES; ++k) {
      tg_mbar_init(bars + k, 2);  // expect_tx arrival (copies) + list arrival
      tg_mbar_init(bars + TGS_STAGES + k, TGS_CONSUMERS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == TGS_CONSUMERS) {
    // ---- producer warp: keeps TGS_STAGES slots in flight (lane 0 issues the bulk copies) and, while the copies
    // fly, writes the slot's band list: the spheres whose depth interval x projected box can meet the chunk's rows,
    // with the tile columns they span.  Consumers only ever look at that list.
    uint64_t stream_once, keep;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(stream_once));
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
    const float fxf = (float)p.fx;
    for (int i = 0, slot = 0, round = 0; i < n; ++i) {
      if (round > 0) tg_mbar_wait(empty0 + 8 * slot, (uint32_t)(round - 1) & 1u, 200);
      const int row0 = kc * R, rows = min(R, H - row0);
      const uint32_t bytes = (uint32_t)rows * (uint32_t)W * 4u;
      unsigned char* dst = tgs_smem + (size_t)slot * TGS_SLOT_BYTES;
      if (lane == 0) {
        tg_mbar_expect_tx(bars + slot, bytes + (uint32_t)sizeof(TgFilterState));
        tg_bulk_g2s(dst, depth + ((size_t)s * H + row0) * W, bytes, bars + slot, stream_once);
        tg_bulk_g2s(dst + TGS_STAGE_BYTES, prep + s, (uint32_t)sizeof(TgFilterState), bars + slot, keep);
      }
      const TgFilterState* ps = prep + s;
      float4* list = reinterpret_cast<float4*>(dst + TGS_STAGE_BYTES + TGS_STATE_BYTES);
      int cnt = 0;
      float ulo = INFINITY, uhi = -INFINITY;
      if ((float)(row0 + rows) >= __ldg(&ps->img_box[1]) && (float)row0 <= __ldg(&ps->img_box[3])) {  // warp-uniform
        const float b0 = ((float)row0 - 0.5f - cyf) * ify, b1 = ((float)(row0 + rows) - 0.5f - cyf) * ify;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int j = lane + 32 * q;
          bool c = false;
          float X = 0.f, Y = 0.f, Z = 0.f;
          if (j < TGM_N_SPHERES) {
            X = __ldg(&ps->cc[3 * j]); Y = __ldg(&ps->cc[3 * j + 1]); Z = __ldg(&ps->cc[3 * j + 2]);
            const float zl = fmaxf(fmaxf(Z - rp, p.zmin - 1e-3f), 0.f), zh = fminf(Z + rp, p.zmax + 1e-3f);
            if (zl <= zh) {
              const float ymin = (b0 >= 0.f ? zl : zh) * b0, ymax = (b1 >= 0.f ? zh : zl) * b1;
              c = ymin <= Y + rp && ymax >= Y - rp;
              if (c) {  // image columns the sphere can cover (as k_filter_prepare's img_box)
                if (zl < 0.01f) {
                  ulo = -INFINITY; uhi = INFINITY;
                } else {
                  const float xa = X - rp, xb = X + rp;
                  ulo = fminf(ulo, cxf + fxf * (xa >= 0.f ? xa / zh : xa / zl) - 1.5f);
                  uhi = fmaxf(uhi, cxf + fxf * (xb >= 0.f ? xb / zl : xb / zh) + 1.5f);
                }
              }
            }
          }
          const uint32_t bq = __ballot_sync(0xffffffffu, c);
          if (c) list[cnt + __popc(bq & ((1u << lane) - 1u))] = make_float4(X, Y, Z, __int_as_float(j));
          cnt += __popc(bq);
        }
#pragma unroll
        for (int of = 16; of; of >>= 1) {
          ulo = fminf(ulo, __shfl_xor_sync(0xffffffffu, ulo, of));
          uhi = fmaxf(uhi, __shfl_xor_sync(0xffffffffu, uhi, of));
        }
      }
      if (lane == 0) {  // float -> int conversion saturates, +-inf are fine
        int4 meta;
        meta.x = cnt;
        meta.y = cnt ? (int)floorf(ulo) : 0x7fffffff;  // pixel columns the band's spheres can cover
        meta.z = cnt ? (int)ceilf(uhi) : -0x7fffffff;
        meta.w = 0;
        *reinterpret_cast<int4*>(dst + TGS_STAGE_BYTES + TGS_STATE_BYTES + TGS_LIST_BYTES) = meta;
      }
      __syncwarp();
      // second arrival on the full barrier (count 2): releases the list and the meta to the consumers
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(full0 + 8 * slot) : "memory");
      s += dS; kc += dK;
      if (kc >= cpf) { kc -= cpf; ++s; }
      if (++slot == TGS_STAGES) { slot = 0; ++round; }
    }
    return;
  }
  // ---- consumers
  // a point can only be inside a sphere of radius r if its camera-frame distance to the centre is below
  // r + 2 mm, and it surely is if that distance is below r - 2 mm: the camera-frame and the world-frame
  // arithmetic differ by rounding only (< 2e-5 m for |p| < 10 m).  The shell between is settled exactly.
  const float rout2 = rp * rp;
  const float rin = fmaxf(radius - 2e-3f, 0.f), rin2 = rin * rin;
  // z > 0 && z >= zmin  <=>  z >= max(zmin, smallest positive float); NaN fails every compare and -inf fails z > 0
  // exactly like np.nan_to_num's 0 / -FLT_MAX; +inf behaves like +FLT_MAX against any zmax
  const uint32_t ok_lo = __float_as_uint(p.zmin > 0.f ? p.zmin : 1.401298464e-45f);
  const uint32_t ok_span = __float_as_uint(p.zmax) - ok_lo;  // the host only takes this path when zmax >= lo
  const int lc = lane & 7, lr = lane >> 3;
  for (int i = 0, slot = 0, round = 0; i < n; ++i) {
    const int row0 = kc * R, rows = min(R, H - row0);
    const unsigned char* base = tgs_smem + (size_t)slot * TGS_SLOT_BYTES;
    const float4* buf = reinterpret_cast<const float4*>(base);
    const TgFilterState& st = *reinterpret_cast<const TgFilterState*>(base + TGS_STAGE_BYTES);
    const float4* list = reinterpret_cast<const float4*>(base + TGS_STAGE_BYTES + TGS_STATE_BYTES);
    uint32_t* mrow = reinterpret_cast<uint32_t*>(mask + ((size_t)s * H + row0) * W);
    tg_mbar_wait(full0 + 8 * slot, (uint32_t)round & 1u, 40);
    const int4 meta = *reinterpret_cast<const int4*>(base + TGS_STAGE_BYTES + TGS_STATE_BYTES + TGS_LIST_BYTES);
    const int cnt = meta.x, u_lo = meta.y, u_hi = meta.z, nq = (cnt + 31) >> 5;
    // one warp = one tile of 32 x (4 TGS_NF4) pixels; lane = TGS_NF4 float4 (4 consecutive pixels each) in rows
    // TH ty + lr (+ 4).  Compact tiles keep the candidate lists short (128 x 1 strips were 25 % slower).
    constexpr int NF = TGS_NF4, TH = 4 * NF, NP = 4 * NF;
    const int nty = (rows + TH - 1) / TH;
    int tx = warp, ty = 0;
    while (tx >= ntx) { tx -= ntx; ++ty; }
    while (ty < nty) {
      const int c4 = 8 * tx + lc, rl = TH * ty + lr;
      const int f = rl * w4 + c4;  // float4 index in the chunk = uint32 index in the chunk's mask rows
      bool have[NF];
      float z[NP];
      uint32_t okb[NF];  // mask bytes of the lane's pixels
#pragma unroll
      for (int h = 0; h < NF; ++h) {
        have[h] = c4 < w4 && rl + 4 * h < rows;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (have[h]) v = buf[f + 4 * h * w4];
        z[4 * h] = v.x; z[4 * h + 1] = v.y; z[4 * h + 2] = v.z; z[4 * h + 3] = v.w;
      }
      // z in [lo, hi] with lo > 0: positive floats order like their bit patterns, everything else (negative, -0, NaN)
      // wraps to a huge unsigned difference -> one subtract + one unsigned compare per pixel
#pragma unroll
      for (int h = 0; h < NF; ++h) {
        okb[h] = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) okb[h] |= (__float_as_uint(z[4 * h + k]) - ok_lo <= ok_span ? 1u : 0u) << (8 * k);
      }
      if (cnt != 0 && 32 * tx + 31 >= u_lo && 32 * tx <= u_hi) {  // warp-uniform: a band sphere reaches these columns
        uint32_t ok = 0;
#pragma unroll
        for (int h = 0; h < NF; ++h)
          ok |= ((okb[h] & 1u) | (okb[h] >> 7 & 2u) | (okb[h] >> 14 & 4u) | (okb[h] >> 21 & 8u)) << (4 * h);
        // depth range of the tile's valid pixels: positive floats order like their bit patterns -> one REDUX each
        uint32_t lo_b = 0x7f800000u, hi_b = 0u;
#pragma unroll
        for (int k = 0; k < NP; ++k)
          if (ok >> k & 1u) { lo_b = min(lo_b, __float_as_uint(z[k])); hi_b = max(hi_b, __float_as_uint(z[k])); }
        const float zlo = __uint_as_float(__reduce_min_sync(0xffffffffu, lo_b));
        const float zhi = __uint_as_float(__reduce_max_sync(0xffffffffu, hi_b));
        uint32_t m[4] = {0u, 0u, 0u, 0u};
        if (zlo <= zhi) {
          // candidate spheres: depth interval meets the tile's depth range and, over that depth range, the tile's
          // back-projected box meets the sphere's box (conservative)
          const int u0 = tx << 5, v0 = row0 + TH * ty;
          const float a0 = ((float)u0 - 0.5f - cxf) * ifx, a1 = ((float)(u0 + 32) - 0.5f - cxf) * ifx;
          const float b0 = ((float)v0 - 0.5f - cyf) * ify, b1 = ((float)(v0 + TH) - 0.5f - cyf) * ify;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (q < nq) {  // warp-uniform
              bool c = false;
              if (lane + 32 * q < cnt) {
                const float4 e = list[lane + 32 * q];
                const float zl = fmaxf(e.z - rp, zlo), zh = fminf(e.z + rp, zhi);
                if (zl <= zh) {
                  const float xmin = (a0 >= 0.f ? zl : zh) * a0, xmax = (a1 >= 0.f ? zh : zl) * a1;
                  const float ymin = (b0 >= 0.f ? zl : zh) * b0, ymax = (b1 >= 0.f ? zh : zl) * b1;
                  c = xmin <= e.x + rp && xmax >= e.x - rp && ymin <= e.y + rp && ymax >= e.y - rp;
                }
              }
              m[q] = __ballot_sync(0xffffffffu, c);
            }
          }
        }
        if ((m[0] | m[1] | m[2] | m[3]) != 0u) {  // warp-uniform: the robot may be visible in this tile
          float ax[4], ay[NF];
#pragma unroll
          for (int k = 0; k < 4; ++k) ax[k] = ((float)(4 * c4 + k) - cxf) * ifx;
#pragma unroll
          for (int h = 0; h < NF; ++h) ay[h] = ((float)(row0 + rl + 4 * h) - cyf) * ify;
          uint32_t inside = 0, shell = 0;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t bits = m[q];
            while (bits) {
              const float4 e = list[32 * q + __ffs(bits) - 1];
              bits &= bits - 1;
#pragma unroll
              for (int k = 0; k < NP; ++k) {
                const float dx = fmaf(z[k], ax[k & 3], -e.x), dy = fmaf(z[k], ay[k >> 2], -e.y), dz = z[k] - e.z;
                const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                inside |= (d2 < rin2 ? 1u : 0u) << k;
                shell |= (d2 < rout2 ? 1u : 0u) << k;
              }
            }
          }
          uint32_t need = shell & ~inside & ok;
          ok &= ~inside;
          while (need) {  // rare (2 mm shells): exact transform + exact test, same arithmetic as the generic kernel
            const int k = __ffs(need) - 1;
            need &= need - 1;
            const int uu = 4 * c4 + (k & 3), vv = row0 + rl + (k & ~3);
            float zz = 0.f;
#pragma unroll
            for (int kk = 0; kk < NP; ++kk)
              if (kk == k) zz = z[kk];  // register select: keeps z[] out of local memory
            if (isinf(zz)) zz = 3.4028234663852886e38f;  // np.nan_to_num (only reachable with zmax = inf)
            const float xc = (float)((((double)(float)uu - p.cx) * (double)zz) / p.fx);
            const float yc = (float)((((double)(float)vv - p.cy) * (double)zz) / p.fy);
            const float wx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[0]), __fmul_rn(yc, st.cam_r[1])), __fmul_rn(zz, st.cam_r[2])), st.cam_t[0]);
            const float wy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[3]), __fmul_rn(yc, st.cam_r[4])), __fmul_rn(zz, st.cam_r[5])), st.cam_t[1]);
            const float wz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, st.cam_r[6]), __fmul_rn(yc, st.cam_r[7])), __fmul_rn(zz, st.cam_r[8])), st.cam_t[2]);
            bool in = false;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint32_t bits = m[q];
              while (bits) {
                const int j = __float_as_int(list[32 * q + __ffs(bits) - 1].w);
                bits &= bits - 1;
                const float dx = __fsub_rn(wx, st.c[3 * j]), dy = __fsub_rn(wy, st.c[3 * j + 1]), dz = __fsub_rn(wz, st.c[3 * j + 2]);
                const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                in |= d2 < p.r2;
              }
            }
            if (in) ok &= ~(1u << k);
          }
#pragma unroll
          for (int h = 0; h < NF; ++h)  // bit k of the nibble lands on bit 8k (k + 7k) of the product, no carries
            okb[h] = ((ok >> (4 * h) & 0xFu) * 0x00204081u) & 0x01010101u;
        }
      }
#pragma unroll
      for (int h = 0; h < NF; ++h)
        if (have[h]) __stcs(mrow + f + 4 * h * w4, okb[h]);
      tx += TGS_CONSUMERS;
      while (tx >= ntx) { tx -= ntx; ++ty; }
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty0 + 8 * slot) : "memory");
    s += dS; kc += dK;
    if (kc >= cpf) { kc -= cpf; ++s; }
    if (++slot == TGS_STAGES) { slot = 0; ++round; }
  }
}

extern "C" int tg_self_filter_depth(const float* depth, int H, int W, int stride, const double* K, const float* states, int64_t S,
                                    float radius, float zmin, float zmax, uint8_t* mask, float* world, void* stream_) {
  TG_CHECK_ARG(H > 0 && W > 0 && stride >= 1 && S >= 0 && S < 65536, "tg_self_filter_depth: bad shape (S < 65536)");
  TG_CHECK_ARG(S == 0 || (depth && K && states && mask), "tg_self_filter_depth: NULL buffer");
  TG_CHECK_ARG(((uintptr_t)depth & 15) == 0, "tg_self_filter_depth: depth must be 16-byte aligned");
  if (S == 0) return TG_OK;
  cudaStream_t st = (cudaStream_t)stream_;
  TgDepthParams p;
  p.H = H; p.W = W; p.stride = stride;
  p.Hs = (H + stride - 1) / stride;
  p.Ws = (W + stride - 1) / stride;
  p.fx = K[0] / stride; p.cx = K[2] / stride;  // depth.py:32-36: K[0,:] /= stride ; K[1,:] /= stride
  p.fy = K[4] / stride; p.cy = K[5] / stride;
  if (stride == 1) { p.fx = K[0]; p.cx = K[2]; p.fy = K[4]; p.cy = K[5]; }
  p.zmin = zmin; p.zmax = zmax;
  p.r2 = (float)((double)radius * (double)radius);
  TgFilterState* prep = nullptr;
  TG_CUDA(cudaMallocAsync(&prep, sizeof(TgFilterState) * (size_t)S, st));
  k_filter_prepare<<<tg_div_up(S, 4), 128, 0, st>>>(states, (int)S, radius, p, prep);
  const int npx = p.Hs * p.Ws;
  if (stride == 1 && world == nullptr && (W & 3) == 0 && ((uintptr_t)mask & 3) == 0 && W * 4 <= TGS_STAGE_BYTES &&
      (long long)S * H < (1ll << 28) && zmax >= (zmin > 0.f ? zmin : 1.401298464e-45f)) {
    static bool attr_set[64] = {};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t smem = (size_t)TGS_STAGES * TGS_SLOT_BYTES + 2 * TGS_STAGES * sizeof(uint64_t);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      TG_CUDA(cudaFuncSetAttribute(k_self_filter_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_set[dev] = true;
    }
    const int R = min(TGS_MAX_ROWS, TGS_STAGE_BYTES / (W * 4));
    const int cpf = tg_div_up(H, R);
    const int total = cpf * (int)S;
    const int blocks = min(sms * TGS_BLOCKS_PER_SM, total);
    k_self_filter_stream<<<blocks, TGS_THREADS, smem, st>>>(depth, prep, p, radius, mask, R, cpf, total);
  } else {
    const int per_state = tg_div_up(npx, TGF_BLOCK * 4);  // many small blocks: measured faster than fewer looping ones
    dim3 grid(per_state, (unsigned)S);
    k_self_filter_depth<4><<<grid, TGF_BLOCK, 0, st>>>(depth, prep, p, radius, mask, world);
  }
  TG_CUDA(cudaGetLastError());
  TG_CUDA(cudaFreeAsync(prep, st));
  return TG_OK;
}
