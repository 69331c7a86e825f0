// ingest.cu — device-resident point-cloud ingest and belief merge (the step directly before K1).
//
// Replaces, with identical results, the numpy chain the reference runs on the host once per scheduler
// iteration (TyGrit/belief_state/pointcloud_scene.py:92-135):
//   depth_to_world_pointcloud   utils/depth.py:13-84      (back-projection f64, world transform f32)
//   crop_sphere                 utils/pointcloud.py:67-80 (XY cylinder, inclusive)
//   filter_ground               utils/pointcloud.py:83-90 (z > threshold)
//   robot self-filter           kinematics/robot_filter.py:13-64
//   voxel_downsample            utils/pointcloud.py:13-31 (first point of every voxel, input order kept)
//   merge_dedup                 utils/pointcloud.py:34-64 (base kept, first new-voxel points of add)
//   points_in_frustum_mask      utils/pointcloud.py:93-133 and the ray-casting carve of
//                               pointcloud_scene.py:218-268
// Every step is a per-point predicate followed by ONE order-preserving stream compaction, so the
// output order equals numpy's boolean-mask order.  HBM-bound: 4 B/pixel in, 12 B/kept point out.
//
// "First point of every voxel" uses an open-addressing hash table keyed by the voxel triple packed
// into 63 bits (21 bits per axis: |coordinate / voxel| < 2^20) with atomicMin on the point index —
// np.unique(axis=0, return_index=True) followed by idx.sort() keeps exactly those points.
#include <math.h>

#include "common.cuh"

#define TGI_BLOCK 256
#define TGI_PER_THREAD 4
#define TGI_TILE (TGI_BLOCK * TGI_PER_THREAD)

// ------------------------------------------------------------------------------------------------
// order-preserving compaction of N x 3 points by a byte flag
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TGI_BLOCK) k_flag_count(const uint8_t* __restrict__ flags, int n, uint32_t* __restrict__ block_count) {
  __shared__ uint32_t warp_sum[TGI_BLOCK / 32];
  const int base = blockIdx.x * TGI_TILE + threadIdx.x * TGI_PER_THREAD;
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < TGI_PER_THREAD; ++k)
    if (base + k < n) c += flags[base + k] != 0;
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < TGI_BLOCK / 32; ++w) t += warp_sum[w];
    block_count[blockIdx.x] = t;
  }
}

// exclusive scan of the block counts in place (single block, any length); total -> *total_out
__global__ void __launch_bounds__(1024) k_scan_counts(uint32_t* __restrict__ v, int nb, int* __restrict__ total_out) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t x = i < nb ? v[i] : 0;
    uint32_t inc = x;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
      const uint32_t w = warp_tot[threadIdx.x];
      uint32_t winc = w;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
        if (threadIdx.x >= o) winc += t;
      }
      warp_tot[threadIdx.x] = winc - w;
    }
    __syncthreads();
    const uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - x;
    if (i < nb) v[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = (int)carry;
}

__global__ void __launch_bounds__(TGI_BLOCK)
k_compact_points(const float* __restrict__ pts, const uint8_t* __restrict__ flags, int n, const uint32_t* __restrict__ block_off,
                 float* __restrict__ out, int cap) {
  __shared__ uint32_t warp_off[TGI_BLOCK / 32];
  const int base = blockIdx.x * TGI_TILE + threadIdx.x * TGI_PER_THREAD;
  bool f[TGI_PER_THREAD];
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < TGI_PER_THREAD; ++k) {
    f[k] = base + k < n && flags[base + k] != 0;
    c += f[k];
  }
  uint32_t inc = c;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((threadIdx.x & 31) >= o) inc += t;
  }
  if ((threadIdx.x & 31) == 31) warp_off[threadIdx.x >> 5] = inc;
  __syncthreads();
  uint32_t pos = block_off[blockIdx.x] + inc - c;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) pos += warp_off[w];
#pragma unroll
  for (int k = 0; k < TGI_PER_THREAD; ++k) {
    if (f[k]) {
      if ((int)pos < cap) {
        const float* s = pts + 3 * (size_t)(base + k);
        float* d = out + 3 * (size_t)pos;
        d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
      }
      ++pos;
    }
  }
}

static int tg_compact_launch(const float* pts, const uint8_t* flags, int64_t n, float* out, int64_t cap, int* count_dev, cudaStream_t st) {
  if (n == 0) {
    TG_CUDA(cudaMemsetAsync(count_dev, 0, sizeof(int), st));
    return TG_OK;
  }
  const int nb = tg_div_up(n, TGI_TILE);
  uint32_t* counts = nullptr;
  TG_CUDA(cudaMallocAsync(&counts, sizeof(uint32_t) * (size_t)nb, st));
  k_flag_count<<<nb, TGI_BLOCK, 0, st>>>(flags, (int)n, counts);
  k_scan_counts<<<1, 1024, 0, st>>>(counts, nb, count_dev);
  k_compact_points<<<nb, TGI_BLOCK, 0, st>>>(pts, flags, (int)n, counts, out, (int)(cap < (1ll << 31) ? cap : (1ll << 31) - 1));
  TG_CUDA(cudaGetLastError());
  TG_CUDA(cudaFreeAsync(counts, st));
  return TG_OK;
}

extern "C" int tg_compact_points(const float* points, const uint8_t* flags, int64_t N, float* out, int64_t capacity, int* count,
                                 void* stream) {
  TG_CHECK_ARG(N >= 0 && N < (1ll << 31) && capacity >= 0 && count != nullptr, "tg_compact_points: bad argument");
  TG_CHECK_ARG(N == 0 || (points && flags && (out || capacity == 0)), "tg_compact_points: NULL buffer");
  return tg_compact_launch(points, flags, N, out, capacity, count, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// first point of every voxel (voxel_downsample / merge_dedup)
// ------------------------------------------------------------------------------------------------
#define TGI_EMPTY 0xFFFFFFFFFFFFFFFFull
#define TGI_UNPACKABLE 0xFFFFFFFFFFFFFFFEull  // coordinate / voxel outside +-2^20 or not finite: never merged

// key of np.floor(points / q).astype(np.int64): the division is evaluated in f32 (f32 array / weak python float)
__device__ __forceinline__ uint64_t tgi_voxel_key(const float* __restrict__ p, float q) {
  const float fx = floorf(__fdiv_rn(p[0], q)), fy = floorf(__fdiv_rn(p[1], q)), fz = floorf(__fdiv_rn(p[2], q));
  const float lim = 1048576.0f;
  if (!(fabsf(fx) < lim && fabsf(fy) < lim && fabsf(fz) < lim)) return TGI_UNPACKABLE;
  const uint64_t ix = (uint64_t)((int)fx + (1 << 20)), iy = (uint64_t)((int)fy + (1 << 20)), iz = (uint64_t)((int)fz + (1 << 20));
  return (ix << 42) | (iy << 21) | iz;
}

__device__ __forceinline__ uint32_t tgi_hash(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (uint32_t)k;
}

// points [0, nb) come from `base`, [nb, nb + na) from `add`; vals[slot] = smallest index with that key
__global__ void __launch_bounds__(256)
k_voxel_insert(const float* __restrict__ base, int nb, const float* __restrict__ add, int na, float q, unsigned long long* __restrict__ keys,
               uint32_t* __restrict__ vals, uint32_t mask, uint32_t* __restrict__ slot_of, int* __restrict__ unpackable) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nb + na) return;
  const float* p = i < nb ? base + 3 * (size_t)i : add + 3 * (size_t)(i - nb);
  const uint64_t key = tgi_voxel_key(p, q);
  if (key == TGI_UNPACKABLE) {
    slot_of[i] = 0xFFFFFFFFu;
    atomicAdd(unpackable, 1);
    return;
  }
  uint32_t slot = tgi_hash(key) & mask;
  for (;;) {
    const unsigned long long old = atomicCAS(&keys[slot], TGI_EMPTY, (unsigned long long)key);
    if (old == TGI_EMPTY || old == key) break;
    slot = (slot + 1) & mask;
  }
  atomicMin(&vals[slot], (uint32_t)i);
  slot_of[i] = slot;
}

__global__ void __launch_bounds__(256)
k_voxel_flags(int nb, int na, const uint32_t* __restrict__ vals, const uint32_t* __restrict__ slot_of, uint8_t* __restrict__ keep_add) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= na) return;
  const uint32_t slot = slot_of[nb + j];
  keep_add[j] = (slot == 0xFFFFFFFFu || vals[slot] == (uint32_t)(nb + j)) ? 1 : 0;
}

extern "C" int tg_voxel_first(const float* base, int64_t nb, const float* add, int64_t na, float voxel, uint8_t* keep_add,
                              int* unpackable, void* stream) {
  TG_CHECK_ARG(nb >= 0 && na >= 0 && nb + na < (1ll << 30), "tg_voxel_first: too many points");
  TG_CHECK_ARG(voxel > 0.f && isfinite(voxel), "tg_voxel_first: voxel size must be positive");
  TG_CHECK_ARG((nb == 0 || base) && (na == 0 || (add && keep_add)), "tg_voxel_first: NULL buffer");
  cudaStream_t st = (cudaStream_t)stream;
  if (unpackable) TG_CUDA(cudaMemsetAsync(unpackable, 0, sizeof(int), st));
  if (na == 0) return TG_OK;
  const int64_t n = nb + na;
  uint32_t size = 1024;
  while ((int64_t)size < 2 * n) size <<= 1;
  unsigned long long* keys = nullptr;
  uint32_t *vals = nullptr, *slot_of = nullptr;
  int* unp = unpackable;
  TG_CUDA(cudaMallocAsync(&keys, sizeof(unsigned long long) * (size_t)size, st));
  TG_CUDA(cudaMallocAsync(&vals, sizeof(uint32_t) * (size_t)size, st));
  TG_CUDA(cudaMallocAsync(&slot_of, sizeof(uint32_t) * (size_t)n, st));
  if (!unp) {
    TG_CUDA(cudaMallocAsync(&unp, sizeof(int), st));
    TG_CUDA(cudaMemsetAsync(unp, 0, sizeof(int), st));
  }
  TG_CUDA(cudaMemsetAsync(keys, 0xFF, sizeof(unsigned long long) * (size_t)size, st));
  TG_CUDA(cudaMemsetAsync(vals, 0xFF, sizeof(uint32_t) * (size_t)size, st));
  k_voxel_insert<<<tg_div_up(n, 256), 256, 0, st>>>(base, (int)nb, add, (int)na, voxel, keys, vals, size - 1, slot_of, unp);
  k_voxel_flags<<<tg_div_up(na, 256), 256, 0, st>>>((int)nb, (int)na, vals, slot_of, keep_add);
  TG_CUDA(cudaGetLastError());
  TG_CUDA(cudaFreeAsync(keys, st));
  TG_CUDA(cudaFreeAsync(vals, st));
  TG_CUDA(cudaFreeAsync(slot_of, st));
  if (!unpackable) TG_CUDA(cudaFreeAsync(unp, st));
  return TG_OK;
}

// ------------------------------------------------------------------------------------------------
// one depth frame -> filtered world cloud (PointCloudScene.update steps 1-4, pointcloud_scene.py:100-121)
// ------------------------------------------------------------------------------------------------
struct TgIngestParams {
  int H, W, stride, Hs, Ws;
  double fx, fy, cx, cy;   // K rows 0 and 1 divided by stride in f64 (depth.py:32-36)
  float rot[9], t[3];      // camera-to-world, cast to f32 (depth.py:79-80)
  float zmin, zmax;
  int crop_on, ground_on, n_centres;
  float crop_cx, crop_cy, crop_r2, ground_z, robot_r2;
};

__global__ void __launch_bounds__(256)
k_ingest_frame(const float* __restrict__ depth, TgIngestParams p, const float* __restrict__ centres, float* __restrict__ world,
               uint8_t* __restrict__ flags) {
  extern __shared__ float sc[];  // robot sphere centres
  for (int j = threadIdx.x; j < 3 * p.n_centres; j += blockDim.x) sc[j] = centres[j];
  __syncthreads();
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= p.Hs * p.Ws) return;
  const int v = pix / p.Ws, u = pix - v * p.Ws;
  float z = __ldg(depth + (size_t)(v * p.stride) * p.W + u * p.stride);
  if (isnan(z)) z = 0.f;  // np.nan_to_num
  else if (isinf(z)) z = z > 0 ? 3.4028234663852886e38f : -3.4028234663852886e38f;
  bool ok = z > 0.f && z >= p.zmin && z <= p.zmax;
  float wx = 0.f, wy = 0.f, wz = 0.f;
  if (ok) {
    const float xc = (float)((((double)(float)u - p.cx) * (double)z) / p.fx);
    const float yc = (float)((((double)(float)v - p.cy) * (double)z) / p.fy);
    wx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, p.rot[0]), __fmul_rn(yc, p.rot[1])), __fmul_rn(z, p.rot[2])), p.t[0]);
    wy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, p.rot[3]), __fmul_rn(yc, p.rot[4])), __fmul_rn(z, p.rot[5])), p.t[1]);
    wz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(xc, p.rot[6]), __fmul_rn(yc, p.rot[7])), __fmul_rn(z, p.rot[8])), p.t[2]);
    if (p.crop_on) {  // crop_sphere: XY cylinder, d2 <= r^2
      const float dx = __fsub_rn(wx, p.crop_cx), dy = __fsub_rn(wy, p.crop_cy);
      ok = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) <= p.crop_r2;
    }
    if (ok && p.ground_on) ok = wz > p.ground_z;
    if (ok && p.n_centres > 0) {
      bool inside = false;
      for (int s = 0; s < p.n_centres; ++s) {
        const float dx = __fsub_rn(wx, sc[3 * s]), dy = __fsub_rn(wy, sc[3 * s + 1]), dz = __fsub_rn(wz, sc[3 * s + 2]);
        inside |= __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)) < p.robot_r2;
      }
      ok = !inside;
    }
  }
  flags[pix] = ok ? 1 : 0;
  float* w = world + 3 * (size_t)pix;
  w[0] = wx; w[1] = wy; w[2] = wz;
}

extern "C" int tg_ingest_depth(const float* depth, int H, int W, int stride, const double* K, const double* T_wc, float zmin,
                               float zmax, float crop_radius, float ground_z, const float* centres, int n_centres,
                               float robot_radius, float* out_points, int64_t capacity, int* count, void* stream) {
  TG_CHECK_ARG(H > 0 && W > 0 && stride >= 1 && depth && K && T_wc && count, "tg_ingest_depth: bad argument");
  TG_CHECK_ARG(n_centres >= 0 && n_centres <= 1024 && (n_centres == 0 || centres), "tg_ingest_depth: at most 1024 robot centres");
  TG_CHECK_ARG(capacity >= 0 && (out_points || capacity == 0), "tg_ingest_depth: NULL output");
  cudaStream_t st = (cudaStream_t)stream;
  TgIngestParams p;
  p.H = H; p.W = W; p.stride = stride;
  p.Hs = (H + stride - 1) / stride;
  p.Ws = (W + stride - 1) / stride;
  p.fx = K[0]; p.cx = K[2]; p.fy = K[4]; p.cy = K[5];
  if (stride > 1) { p.fx = K[0] / stride; p.cx = K[2] / stride; p.fy = K[4] / stride; p.cy = K[5] / stride; }
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) p.rot[3 * r + c] = (float)T_wc[4 * r + c];
    p.t[r] = (float)T_wc[4 * r + 3];
  }
  p.zmin = zmin; p.zmax = zmax;
  p.crop_on = crop_radius >= 0.f;
  p.crop_cx = p.t[0]; p.crop_cy = p.t[1];  // pointcloud_scene.py:110-111: centre = camera position
  p.crop_r2 = (float)((double)crop_radius * (double)crop_radius);
  p.ground_on = !isnan(ground_z);
  p.ground_z = ground_z;
  p.n_centres = n_centres;
  p.robot_r2 = (float)((double)robot_radius * (double)robot_radius);
  const int npx = p.Hs * p.Ws;
  float* world = nullptr;
  uint8_t* flags = nullptr;
  TG_CUDA(cudaMallocAsync(&world, sizeof(float) * 3 * (size_t)npx, st));
  TG_CUDA(cudaMallocAsync(&flags, (size_t)npx, st));
  k_ingest_frame<<<tg_div_up(npx, 256), 256, sizeof(float) * 3 * (size_t)n_centres, st>>>(depth, p, centres, world, flags);
  TG_CUDA(cudaGetLastError());
  const int rc = tg_compact_launch(world, flags, npx, out_points, capacity, count, st);
  TG_CUDA(cudaFreeAsync(world, st));
  TG_CUDA(cudaFreeAsync(flags, st));
  return rc;
}

// ------------------------------------------------------------------------------------------------
// carving the accumulated map: frustum mask (utils/pointcloud.py:93-133) and free-space ray test
// (pointcloud_scene.py:218-262).  keep[i] = 0 for points the reference removes.
// ------------------------------------------------------------------------------------------------
struct TgCarveParams {
  float rot[9], t[3];
  double fx, fy, cx, cy;
  float zmin, zmax, merge_radius;
  int width, height, mode;  // mode 0: frustum, 1: ray casting against `depth` (height x width)
};

__global__ void __launch_bounds__(256)
k_carve_flags(const float* __restrict__ pts, int n, TgCarveParams p, const float* __restrict__ depth, uint8_t* __restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float dx = __fsub_rn(pts[3 * (size_t)i], p.t[0]), dy = __fsub_rn(pts[3 * (size_t)i + 1], p.t[1]),
              dz = __fsub_rn(pts[3 * (size_t)i + 2], p.t[2]);
  // pc = (points - t) @ rot : pc[k] = sum_i d[i] * rot[i][k]  (f32)
  const float pcx = __fadd_rn(__fadd_rn(__fmul_rn(dx, p.rot[0]), __fmul_rn(dy, p.rot[3])), __fmul_rn(dz, p.rot[6]));
  const float pcy = __fadd_rn(__fadd_rn(__fmul_rn(dx, p.rot[1]), __fmul_rn(dy, p.rot[4])), __fmul_rn(dz, p.rot[7]));
  const float pcz = __fadd_rn(__fadd_rn(__fmul_rn(dx, p.rot[2]), __fmul_rn(dy, p.rot[5])), __fmul_rn(dz, p.rot[8]));
  bool remove = false;
  if (p.mode == 0) {
    if (pcz >= p.zmin && pcz <= p.zmax) {
      const double u = ((double)pcx * p.fx) / (double)pcz + p.cx, v = ((double)pcy * p.fy) / (double)pcz + p.cy;
      remove = u >= 0.0 && u < (double)p.width && v >= 0.0 && v < (double)p.height;
    }
  } else if (pcz > 1e-3f) {
    const double u = ((double)pcx * p.fx) / (double)pcz + p.cx, v = ((double)pcy * p.fy) / (double)pcz + p.cy;
    const double ur = rint(u), vr = rint(v);
    if (ur >= 0.0 && ur < (double)p.width && vr >= 0.0 && vr < (double)p.height) {
      const float zm = __ldg(depth + (size_t)(int)vr * p.width + (int)ur);
      remove = zm > 0.f && isfinite(zm) && pcz < __fsub_rn(zm, p.merge_radius);
    }
  }
  keep[i] = remove ? 0 : 1;
}

extern "C" int tg_carve_flags(const float* points, int64_t N, int mode, const double* K, const double* T_wc, float zmin, float zmax,
                              int width, int height, const float* depth, float merge_radius, uint8_t* keep, void* stream) {
  TG_CHECK_ARG(N >= 0 && N < (1ll << 31) && (mode == 0 || mode == 1) && K && T_wc, "tg_carve_flags: bad argument");
  TG_CHECK_ARG(N == 0 || (points && keep), "tg_carve_flags: NULL buffer");
  TG_CHECK_ARG(mode == 0 || depth, "tg_carve_flags: ray casting needs the depth frame");
  if (N == 0) return TG_OK;
  TgCarveParams p;
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) p.rot[3 * r + c] = (float)T_wc[4 * r + c];
    p.t[r] = (float)T_wc[4 * r + 3];
  }
  p.fx = K[0]; p.cx = K[2]; p.fy = K[4]; p.cy = K[5];
  p.zmin = zmin; p.zmax = zmax; p.merge_radius = merge_radius;
  p.width = width; p.height = height; p.mode = mode;
  k_carve_flags<<<tg_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(points, (int)N, p, depth, keep);
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}
