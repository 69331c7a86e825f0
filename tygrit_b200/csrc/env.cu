// env.cu — K1: point cloud -> sorted cell table + distance-bound grid, all in HBM.
//
// Replaces the CAPT rebuild behind VampPreviewPlanner.update_environment / add_pointcloud
// (TyGrit/planning/fetch/vamp_preview.py:183-240), which the reference pays on every scheduler
// iteration through a Python list-of-lists (pts.tolist(), :222-224).
//
// Pipeline on the caller's stream (HBM-bound, 28 B/point compulsory + 4 B/voxel for the grid):
//   k_bounds      min/max of the finite points (skipped with fixed bounds)
//   k_count       per point: cell key, rank inside its cell (global atomic), mark its voxel
//   k_scan_*      exclusive scan of the cell histogram -> cell_start
//   k_scatter     counting-sort scatter into float4 points (x fastest key => rows are contiguous)
//   k_edt_xy/z    exact truncated Euclidean distance transform between voxel centres from an
//                 occupancy bitmask, shared-memory tiles + DPX add-min, one byte per voxel
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "fetch_model.cuh"

// ------------------------------------------------------------------------------------------------
// error text
// ------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void tg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* tg_last_error(void) { return g_err; }
extern "C" int tg_version(void) { return 100; }

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int f2ord(float f) {  // monotone float -> int
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
static inline float ord2f(int i) {
  int j = i >= 0 ? i : i ^ 0x7FFFFFFF;
  float f;
  memcpy(&f, &j, 4);
  return f;
}

__global__ void k_bounds_init(int* b) {
  if (threadIdx.x < 3) b[threadIdx.x] = 0x7FFFFFFF;
  else if (threadIdx.x < 6) b[threadIdx.x] = (int)0x80000000;
  else if (threadIdx.x < 8) b[threadIdx.x] = 0;
}

__global__ void __launch_bounds__(256) k_bounds(const float* __restrict__ xyz, int n, int* __restrict__ b) {
  int lo[3] = {0x7FFFFFFF, 0x7FFFFFFF, 0x7FFFFFFF};
  int hi[3] = {(int)0x80000000, (int)0x80000000, (int)0x80000000};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
      lo[0] = min(lo[0], f2ord(x)); hi[0] = max(hi[0], f2ord(x));
      lo[1] = min(lo[1], f2ord(y)); hi[1] = max(hi[1], f2ord(y));
      lo[2] = min(lo[2], f2ord(z)); hi[2] = max(hi[2], f2ord(z));
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o; o >>= 1) {
      lo[a] = min(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = max(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(&b[a], lo[a]);
      atomicMax(&b[3 + a], hi[a]);
    }
  }
}

// per point: cell key + rank in cell; mark the df voxel.  Points that are not finite or fall
// outside the grids are dropped (key = 0xFFFFFFFF) and counted.
__global__ void __launch_bounds__(256)
k_count(const float* __restrict__ xyz, int n, TgGrid g, uint32_t* __restrict__ occ_bits, int nwx, uint32_t* __restrict__ cbits, int cnwx,
        uint32_t* __restrict__ cell_count, uint32_t* __restrict__ key, uint32_t* __restrict__ rank, int* __restrict__ dropped) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  uint32_t k = 0xFFFFFFFFu;
  if (isfinite(x) && isfinite(y) && isfinite(z)) {
    // df voxel: same arithmetic as the query side (one FMA-free multiply + magic add)
    const float fx = fmaf(x, g.vpm, g.bias[0]), fy = fmaf(y, g.vpm, g.bias[1]), fz = fmaf(z, g.vpm, g.bias[2]);
    const int cx = (int)floorf(x * g.cpmx) - g.cm[0];
    const int cy = (int)floorf(y * g.cpm) - g.cm[1];
    const int cz = (int)floorf(z * g.cpm) - g.cm[2];
    const bool in_df = fx >= TG_MAGIC && fx <= g.hi[0] && fy >= TG_MAGIC && fy <= g.hi[1] && fz >= TG_MAGIC && fz <= g.hi[2];
    const bool in_cells = cx >= 0 && cx < g.cnx && cy >= 0 && cy < g.cny && cz >= 0 && cz < g.cnz;
    if (in_df && in_cells) {
      const unsigned ix = __float_as_uint(fx) - TG_MAGIC_BITS, iy = __float_as_uint(fy) - TG_MAGIC_BITS,
                     iz = __float_as_uint(fz) - TG_MAGIC_BITS;
      // test before set: bits only ever get set, so a stale read can only cost a redundant atomic (the coarse words are
      // hit by hundreds of points each)
      uint32_t* w = &occ_bits[((size_t)iz * g.ny + iy) * nwx + (ix >> 5)];
      if (!((__ldcg(w) >> (ix & 31)) & 1u)) atomicOr(w, 1u << (ix & 31));
      const unsigned jx = ix / (unsigned)g.cg_r, jy = iy / (unsigned)g.cg_r, jz = iz / (unsigned)g.cg_r;
      uint32_t* cw = &cbits[(jz * g.gny + jy) * cnwx + (jx >> 5)];
      if (!((__ldcg(cw) >> (jx & 31)) & 1u)) atomicOr(cw, 1u << (jx & 31));
      k = (uint32_t)((cz * g.cny + cy) * g.cnx + cx);
      rank[i] = atomicAdd(&cell_count[k], 1u);
    }
  }
  key[i] = k;
  if (k == 0xFFFFFFFFu) atomicAdd(dropped, 1);
}

// ---- exclusive scan over the cell histogram (3 kernels, 2048 items per block) -------------------
#define SCAN_ITEMS 2048
__global__ void __launch_bounds__(256) k_scan_block_sums(const uint32_t* __restrict__ in, int n, uint32_t* __restrict__ sums) {
  __shared__ uint32_t warp_sum[8];
  const int base = blockIdx.x * SCAN_ITEMS;
  uint32_t s = 0;
  for (int j = threadIdx.x; j < SCAN_ITEMS; j += 256)
    if (base + j < n) s += in[base + j];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += warp_sum[w];
    sums[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(1024) k_scan_sums(uint32_t* sums, int nb) {  // single block, in place, exclusive
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    uint32_t v = i < nb ? sums[i] : 0, inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = warp_tot[threadIdx.x], winc = w;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
        if (threadIdx.x >= o) winc += t;
      }
      warp_tot[threadIdx.x] = winc - w;  // exclusive warp offsets
    }
    __syncthreads();
    const uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - v;
    if (i < nb) sums[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
}
// out[i] = exclusive prefix; out[n] = total.  Block re-scans its 2048 items serially per thread (8 each).
__global__ void __launch_bounds__(256)
k_scan_apply(const uint32_t* __restrict__ in, int n, const uint32_t* __restrict__ sums, uint32_t* __restrict__ out) {
  __shared__ uint32_t warp_tot[8];
  const int base = blockIdx.x * SCAN_ITEMS + threadIdx.x * 8;
  uint32_t v[8], s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    v[j] = (base + j < n) ? in[base + j] : 0;
    s += v[j];
  }
  uint32_t inc = s;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((threadIdx.x & 31) >= o) inc += t;
  }
  if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
  __syncthreads();
  uint32_t off = sums[blockIdx.x];
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) off += warp_tot[w];
  uint32_t run = off + inc - s;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (base + j < n) out[base + j] = run;
    run += v[j];
    if (base + j == n - 1) out[n] = run;
  }
}

__global__ void __launch_bounds__(256)
k_scatter(const float* __restrict__ xyz, int n, const uint32_t* __restrict__ key, const uint32_t* __restrict__ rank,
          const uint32_t* __restrict__ cell_start, float4* __restrict__ pts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t k = key[i];
  if (k == 0xFFFFFFFFu) return;
  pts[cell_start[k] + rank[i]] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0.f);
}

// ---- shared-memory staged radix sort of the points by cell key (the default; -DTG_SORT_GLOBAL_ATOMICS = the
// round-1 pipeline above: one global atomic per point + a 3-kernel scan over all cells) -------------------------------
// Two-level MSD counting sort.  A BUCKET is a run of 2^cpb_shift consecutive cells (a slab of cell rows):
//   A  k_sort_hist       per block of 2048 points: voxel / coarse occupancy bits, cell key, bucket histogram in SHARED
//                        memory, then ONE global atomic per (block, non-empty bucket)
//   B  k_sort_scan       exclusive scan of the bucket counts (a single block; <= a few thousand buckets)
//   C  k_sort_partition  per block: shared-memory histogram again, one reservation per (block, bucket), ranks from
//                        shared-memory atomics, points written as float4 (x, y, z, key) grouped by bucket
//   D  k_sort_bucket     one block per bucket: histogram of the bucket's cells in shared memory, block scan ->
//                        cell_start for its cells (no global scan over all cells), second sweep places the points
#define TGS_CPB_MAX 8192      // cells per bucket: 2048 (8 KB of shared-memory counters in pass D), doubled while the bucket count
                              // would exceed TGS_MAX_BUCKETS; a power of two (passed as a shift)
#define TGS_TILE 2048         // points per block in passes A and C
#define TGS_MAX_BUCKETS 6144  // shared-memory histograms of passes A and C (24 KB / 48 KB)

__device__ __forceinline__ uint32_t tgs_key(const TgGrid& g, float x, float y, float z, bool mark, uint32_t* occ_bits, int nwx, uint32_t* cbits,
                                            int cnwx) {
  if (!(isfinite(x) && isfinite(y) && isfinite(z))) return 0xFFFFFFFFu;
  const float fx = fmaf(x, g.vpm, g.bias[0]), fy = fmaf(y, g.vpm, g.bias[1]), fz = fmaf(z, g.vpm, g.bias[2]);
  const int cx = (int)floorf(x * g.cpmx) - g.cm[0], cy = (int)floorf(y * g.cpm) - g.cm[1], cz = (int)floorf(z * g.cpm) - g.cm[2];
  const bool in_df = fx >= TG_MAGIC && fx <= g.hi[0] && fy >= TG_MAGIC && fy <= g.hi[1] && fz >= TG_MAGIC && fz <= g.hi[2];
  const bool in_cells = cx >= 0 && cx < g.cnx && cy >= 0 && cy < g.cny && cz >= 0 && cz < g.cnz;
  if (!(in_df && in_cells)) return 0xFFFFFFFFu;
  if (mark) {
    const unsigned ix = __float_as_uint(fx) - TG_MAGIC_BITS, iy = __float_as_uint(fy) - TG_MAGIC_BITS, iz = __float_as_uint(fz) - TG_MAGIC_BITS;
    uint32_t* w = &occ_bits[((size_t)iz * g.ny + iy) * nwx + (ix >> 5)];
    if (!((__ldcg(w) >> (ix & 31)) & 1u)) atomicOr(w, 1u << (ix & 31));  // test before set: bits only ever get set
    const unsigned jx = ix / (unsigned)g.cg_r, jy = iy / (unsigned)g.cg_r, jz = iz / (unsigned)g.cg_r;
    uint32_t* cw = &cbits[(jz * g.gny + jy) * cnwx + (jx >> 5)];
    if (!((__ldcg(cw) >> (jx & 31)) & 1u)) atomicOr(cw, 1u << (jx & 31));
  }
  return (uint32_t)((cz * g.cny + cy) * g.cnx + cx);
}

__global__ void __launch_bounds__(256)
k_sort_hist(const float* __restrict__ xyz, int n, TgGrid g, uint32_t* __restrict__ occ_bits, int nwx, uint32_t* __restrict__ cbits, int cnwx,
            int nbuckets, int cpb_shift, uint32_t* __restrict__ key, uint32_t* __restrict__ bucket_count, int* __restrict__ dropped) {
  extern __shared__ uint32_t s_hist[];
  for (int b = threadIdx.x; b < nbuckets; b += 256) s_hist[b] = 0;
  __syncthreads();
  int ndrop = 0;
  for (int j = threadIdx.x; j < TGS_TILE; j += 256) {
    const int i = blockIdx.x * TGS_TILE + j;
    if (i >= n) break;
    const uint32_t k = tgs_key(g, xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2], true, occ_bits, nwx, cbits, cnwx);
    key[i] = k;
    if (k == 0xFFFFFFFFu) ++ndrop;
    else atomicAdd(&s_hist[k >> cpb_shift], 1u);
  }
  if (ndrop) atomicAdd(dropped, ndrop);
  __syncthreads();
  for (int b = threadIdx.x; b < nbuckets; b += 256)
    if (s_hist[b]) atomicAdd(&bucket_count[b], s_hist[b]);
}

// exclusive scan of the bucket counts: start[b], start[nbuckets] = kept points; cursor[b] = start[b]
__global__ void __launch_bounds__(1024) k_sort_scan(const uint32_t* __restrict__ count, int nb, uint32_t* __restrict__ start, uint32_t* __restrict__ cursor) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < nb ? count[i] : 0;
    uint32_t inc = v;
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
    const uint32_t excl = carry + warp_tot[threadIdx.x >> 5] + inc - v;
    if (i < nb) {
      start[i] = excl;
      cursor[i] = excl;
    }
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) start[nb] = carry;
}

__global__ void __launch_bounds__(256)
k_sort_partition(const float* __restrict__ xyz, int n, const uint32_t* __restrict__ key, int nbuckets, int cpb_shift,
                 uint32_t* __restrict__ cursor, float4* __restrict__ tmp) {
  extern __shared__ uint32_t s_mem[];  // [nbuckets] counts -> base offsets, [nbuckets] running ranks
  uint32_t* s_base = s_mem;
  uint32_t* s_rank = s_mem + nbuckets;
  for (int b = threadIdx.x; b < nbuckets; b += 256) s_base[b] = s_rank[b] = 0;
  __syncthreads();
  for (int j = threadIdx.x; j < TGS_TILE; j += 256) {
    const int i = blockIdx.x * TGS_TILE + j;
    if (i >= n) break;
    const uint32_t k = key[i];
    if (k != 0xFFFFFFFFu) atomicAdd(&s_base[k >> cpb_shift], 1u);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nbuckets; b += 256)
    if (s_base[b]) s_base[b] = atomicAdd(&cursor[b], s_base[b]);  // one reservation per (block, bucket)
  __syncthreads();
  for (int j = threadIdx.x; j < TGS_TILE; j += 256) {
    const int i = blockIdx.x * TGS_TILE + j;
    if (i >= n) break;
    const uint32_t k = key[i];
    if (k == 0xFFFFFFFFu) continue;
    const uint32_t b = k >> cpb_shift;
    const uint32_t dst = s_base[b] + atomicAdd(&s_rank[b], 1u);
    tmp[dst] = make_float4(xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2], __uint_as_float(k));
  }
}

__global__ void __launch_bounds__(256)
k_sort_bucket(const float4* __restrict__ tmp, const uint32_t* __restrict__ bucket_start, int nbuckets, int cpb_shift, int ncell,
              uint32_t* __restrict__ cell_start, float4* __restrict__ pts) {
  extern __shared__ uint32_t s_cnt[];  // [cpb]
  __shared__ uint32_t s_warp[8];
  const int b = blockIdx.x, cpb = 1 << cpb_shift;
  const uint32_t p0 = bucket_start[b], p1 = bucket_start[b + 1];
  const uint32_t c0 = (uint32_t)b << cpb_shift;
  const int nc = min(cpb, ncell - (int)c0);
  for (int c = threadIdx.x; c < cpb; c += 256) s_cnt[c] = 0;
  __syncthreads();
  for (uint32_t p = p0 + threadIdx.x; p < p1; p += 256) atomicAdd(&s_cnt[__float_as_uint(__ldg(&tmp[p].w)) - c0], 1u);
  __syncthreads();
  // block exclusive scan of the cpb counters: cpb / 256 consecutive counters per thread
  const int per = cpb >> 8;
  uint32_t sum = 0;
  for (int j = 0; j < per; ++j) sum += s_cnt[threadIdx.x * per + j];
  uint32_t inc = sum;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((threadIdx.x & 31) >= o) inc += t;
  }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = inc;
  __syncthreads();
  uint32_t run = p0 + inc - sum;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) run += s_warp[w];
  for (int j = 0; j < per; ++j) {
    const int c = threadIdx.x * per + j;
    const uint32_t v = s_cnt[c];
    s_cnt[c] = run;  // becomes the cell's write cursor
    if (c < nc) cell_start[c0 + c] = run;
    run += v;
  }
  if (b == nbuckets - 1 && threadIdx.x == 0) cell_start[ncell] = p1;
  __syncthreads();
  for (uint32_t p = p0 + threadIdx.x; p < p1; p += 256) {
    const float4 v = __ldg(&tmp[p]);
    const uint32_t pos = atomicAdd(&s_cnt[__float_as_uint(v.w) - c0], 1u);
    pts[pos] = make_float4(v.x, v.y, v.z, 0.f);
  }
}

// ---- truncated exact EDT --------------------------------------------------------------------------
// Exact Euclidean distance transform between voxel centres, truncated at W voxels, separable:
//   X  nearest occupied voxel in the same x-row, straight from the occupancy BITMASK (funnel shift
//      + clz/ffs: O(1) per voxel, no scan);
//   Y  g2(y) = min_k gX(y+k)^2 + k^2   (fused with X: gX is never written to HBM);
//   Z  d2(z) = min_k g2(z+k)  + k^2,  code = d2 >= W^2 ? 255 : floor(sqrt(d2) * 255 / W).
// The min-plus taps run on packed u16x2 pairs (two adjacent x per thread) with the DPX instruction
// VIADDMNMX.U16x2 (__viaddmin_u16x2): one shared-memory load + one instruction per tap per 2 voxels.
// Squared distances stay below 2*31^2; TGE_SENT marks "nothing within W" and survives the adds.
#define TGE_SENT 0x7FFFu
#define TGE_SENT2 0x7FFF7FFFu
#define TGE_SEG 96     // voxels along the pass axis per block (plus a halo of W on both sides)
#define TGE_THREADS 256

__device__ __forceinline__ unsigned tge_row_dist2(const uint32_t* __restrict__ row, int nwx, int x, int W) {
  const int wi = x >> 5, sh = x & 31;
  const uint32_t wm = wi > 0 ? __ldg(row + wi - 1) : 0u;
  const uint32_t w0 = wi < nwx ? __ldg(row + wi) : 0u;
  const uint32_t wp = wi + 1 < nwx ? __ldg(row + wi + 1) : 0u;
  const uint32_t r32 = __funnelshift_r(w0, wp, sh);       // bit 0 = voxel x, bit j = x + j
  const uint32_t l32 = __funnelshift_l(wm, w0, 31 - sh);  // bit 31 = voxel x, bit 31 - j = x - j
  const int dr = r32 ? __ffs(r32) - 1 : 64;
  const int dl = l32 ? __clz(l32) : 64;
  const int d = min(dl, dr);
  return d <= W ? (unsigned)(d * d) : TGE_SENT;
}

// Register-blocked taps: one thread produces TGE_T consecutive outputs along the pass axis from TGE_T + 2 WB loaded
// rows (one LDS per ~10 taps instead of one per tap: the passes were shared-memory-bandwidth bound).  WB >= W is the
// window rounded up to a multiple of 4 so that the tap offsets and k^2 are compile-time constants with 7
// instantiations; the extra taps only add candidates > W^2, which the final quantisation maps to 255 like "nothing
// within W", so the codes are unchanged (bit-exact test against scipy's EDT).
#define TGE_T 12
template <int WB>
__device__ __forceinline__ void tge_taps_blocked(const uint32_t* __restrict__ col /* tile + r0 * 32 + lane */, unsigned (&best)[TGE_T]) {
#pragma unroll
  for (int t = 0; t < TGE_T; ++t) best[t] = TGE_SENT2;
#pragma unroll
  for (int j = -WB; j < TGE_T + WB; ++j) {
    const unsigned v = col[32 * j];
#pragma unroll
    for (int t = 0; t < TGE_T; ++t) {
      const int k = j - t;
      if (k >= -WB && k <= WB) best[t] = __viaddmin_u16x2(v, (unsigned)(k * k) * 0x10001u, best[t]);
    }
  }
#pragma unroll
  for (int t = 0; t < TGE_T; ++t) best[t] = __vminu2(best[t], TGE_SENT2);
}

// passes X + Y.  grid (ceil(nx/64), ceil(ny/TGE_SEG), nz); g2 is u16 [nz][ny][nxp], nxp even.
template <int WB>
__global__ void __launch_bounds__(TGE_THREADS)
k_edt_xy(const uint32_t* __restrict__ bits, uint32_t* __restrict__ g2w, int nx, int nxp, int nwx, int ny, int W) {
  extern __shared__ uint32_t tile[];  // [TGE_SEG + 2 WB][32] packed pairs
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = blockIdx.x * 64 + 2 * lane, y0 = blockIdx.y * TGE_SEG, z = blockIdx.z;
  for (int r = warp; r < TGE_SEG + 2 * WB; r += TGE_THREADS / 32) {
    const int y = y0 - WB + r;
    unsigned v = TGE_SENT2;
    if (y >= 0 && y < ny && x < nx) {
      const uint32_t* row = bits + ((size_t)z * ny + y) * nwx;
      const unsigned a = tge_row_dist2(row, nwx, x, W);
      const unsigned b = (x + 1 < nx) ? tge_row_dist2(row, nwx, x + 1, W) : TGE_SENT;
      v = a | (b << 16);
    }
    tile[r * 32 + lane] = v;
  }
  __syncthreads();
  for (int r0 = TGE_T * warp; r0 < TGE_SEG; r0 += TGE_T * (TGE_THREADS / 32)) {
    if (y0 + r0 >= ny) break;  // warp-uniform
    // nothing within W in any of the rows this strip of outputs can see (most of a room's volume): the taps would only
    // reproduce the sentinel.  One LDS + AND per input row instead of ~10 DPX taps per row.
    const uint32_t* col = tile + (WB + r0) * 32 + lane;
    unsigned all = TGE_SENT2;
#pragma unroll 4
    for (int j = -WB; j < TGE_T + WB; ++j) all &= col[32 * j];
    unsigned best[TGE_T];
    if (__all_sync(0xffffffffu, all == TGE_SENT2)) {
#pragma unroll
      for (int t = 0; t < TGE_T; ++t) best[t] = TGE_SENT2;
    } else {
      tge_taps_blocked<WB>(col, best);
    }
    if (x < nxp) {
#pragma unroll
      for (int t = 0; t < TGE_T; ++t) {
        const int y = y0 + r0 + t;
        if (y < ny) g2w[(((size_t)z * ny + y) * nxp + x) >> 1] = best[t];
      }
    }
  }
}

// pass Z + quantisation.  grid (ceil(nx/64), ny, ceil(nz/TGE_SEG)).
template <int WB>
__global__ void __launch_bounds__(TGE_THREADS)
k_edt_z(const uint32_t* __restrict__ g2w, uint8_t* __restrict__ df, cudaSurfaceObject_t surf, int nx, int nxp, int ny, int nz, int W) {
  extern __shared__ uint32_t tile[];  // [TGE_SEG + 2 WB][32] packed pairs
  __shared__ uint8_t lut[1024];       // d2 -> code for d2 < W^2 (W <= 31)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = blockIdx.x * 64 + 2 * lane, y = blockIdx.y, z0 = blockIdx.z * TGE_SEG;
  const float scale = 255.0f / (float)W;
  for (int d2 = threadIdx.x; d2 < 1024; d2 += TGE_THREADS)
    lut[d2] = d2 < W * W ? (uint8_t)min(254, (int)floorf(sqrtf((float)d2) * scale)) : (uint8_t)255;
  for (int r = warp; r < TGE_SEG + 2 * WB; r += TGE_THREADS / 32) {
    const int z = z0 - WB + r;
    unsigned v = TGE_SENT2;
    if (z >= 0 && z < nz && x < nxp) v = __ldg(g2w + ((((size_t)z * ny + y) * nxp + x) >> 1));
    tile[r * 32 + lane] = v;
  }
  __syncthreads();
  for (int r0 = TGE_T * warp; r0 < TGE_SEG; r0 += TGE_T * (TGE_THREADS / 32)) {
    if (z0 + r0 >= nz) break;  // warp-uniform
    const uint32_t* col = tile + (WB + r0) * 32 + lane;
    unsigned all = TGE_SENT2;  // see k_edt_xy: skip the taps when no input of this strip is below the sentinel
#pragma unroll 4
    for (int j = -WB; j < TGE_T + WB; ++j) all &= col[32 * j];
    unsigned best[TGE_T];
    if (__all_sync(0xffffffffu, all == TGE_SENT2)) {
#pragma unroll
      for (int t = 0; t < TGE_T; ++t) best[t] = TGE_SENT2;
    } else {
      tge_taps_blocked<WB>(col, best);
    }
    if (x < nx) {
#pragma unroll
      for (int t = 0; t < TGE_T; ++t) {
        const int z = z0 + r0 + t;
        if (z < nz) {
          const unsigned lo = best[t] & 0xFFFFu, hi = best[t] >> 16;
          uint8_t* out = df + ((size_t)z * ny + y) * nx + x;
          const uint8_t c0 = lo < 1024u ? lut[lo] : (uint8_t)255, c1 = hi < 1024u ? lut[hi] : (uint8_t)255;
          out[0] = c0;
          if (x + 1 < nx) out[1] = c1;
          // the block-linear copy phase 1 reads through the texture path is written here too (no separate 3-D copy)
          if (surf) {
            surf3Dwrite(c0, surf, x, y, z);
            if (x + 1 < nx) surf3Dwrite(c1, surf, x + 1, y, z);
          }
        }
      }
    }
  }
}

// pass Z of the COARSE far-field grid: the same transform on blocks of cg_r^3 voxels (a block is occupied when any of
// its voxels is); the output code is a LOWER bound in units of cq metres on the distance from ANY point of the block to
// the cloud:  lo = |centre - nearest occupied block centre| - 2 * (half diagonal of a block)  (both points can sit
// anywhere in their blocks), floored.  Nothing within W blocks: lo from W.
template <int WB>
__global__ void __launch_bounds__(TGE_THREADS)
k_edt_z_coarse(const uint32_t* __restrict__ g2w, uint8_t* __restrict__ cg, int nx, int nxp, int ny, int nz, int W, float cell, float cq) {
  extern __shared__ uint32_t tile[];
  __shared__ uint8_t lut[1024];
  __shared__ uint8_t sat;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = blockIdx.x * 64 + 2 * lane, y = blockIdx.y, z0 = blockIdx.z * TGE_SEG;
  const float two_h = 1.7320509f * cell + 3e-5f;
  for (int d2 = threadIdx.x; d2 <= 1024; d2 += TGE_THREADS) {
    const float d = d2 < W * W ? sqrtf((float)d2) : (float)W;
    const uint8_t c = (uint8_t)min(255, max(0, (int)floorf((d * cell * 0.999999f - two_h) / cq)));
    if (d2 < 1024) lut[d2] = c;
    if (d2 == 1024) sat = c;
  }
  for (int r = warp; r < TGE_SEG + 2 * WB; r += TGE_THREADS / 32) {
    const int z = z0 - WB + r;
    unsigned v = TGE_SENT2;
    if (z >= 0 && z < nz && x < nxp) v = __ldg(g2w + ((((size_t)z * ny + y) * nxp + x) >> 1));
    tile[r * 32 + lane] = v;
  }
  __syncthreads();
  if (x >= nx) return;
  for (int r0 = TGE_T * warp; r0 < TGE_SEG; r0 += TGE_T * (TGE_THREADS / 32)) {
    if (z0 + r0 >= nz) break;
    unsigned best[TGE_T];
    tge_taps_blocked<WB>(tile + (WB + r0) * 32 + lane, best);
#pragma unroll
    for (int t = 0; t < TGE_T; ++t) {
      const int z = z0 + r0 + t;
      if (z < nz) {
        const unsigned lo = best[t] & 0xFFFFu, hi = best[t] >> 16;
        uint8_t* out = cg + ((size_t)z * ny + y) * nx + x;
        out[0] = lo < 1024u ? lut[lo] : sat;
        if (x + 1 < nx) out[1] = hi < 1024u ? lut[hi] : sat;
      }
    }
  }
}

template <int WB>
static void tg_launch_edt_coarse(const uint32_t* cbits, uint32_t* g2c, uint8_t* cg, const TgGrid& g, int W, float cell, cudaStream_t st) {
  const int nxp = (g.gnx + 1) & ~1, nwx = (g.gnx + 31) / 32;
  const size_t tile_bytes = (size_t)(TGE_SEG + 2 * WB) * 32 * sizeof(uint32_t);
  k_edt_xy<WB><<<dim3(tg_div_up(nxp, 64), tg_div_up(g.gny, TGE_SEG), g.gnz), TGE_THREADS, tile_bytes, st>>>(cbits, g2c, g.gnx, nxp, nwx, g.gny, W);
  k_edt_z_coarse<WB><<<dim3(tg_div_up(g.gnx, 64), g.gny, tg_div_up(g.gnz, TGE_SEG)), TGE_THREADS, tile_bytes, st>>>(g2c, cg, g.gnx, nxp, g.gny, g.gnz, W,
                                                                                                               cell, g.cg_q);
}

template <int WB>
static void tg_launch_edt(const uint32_t* occ_bits, uint32_t* g2, uint8_t* df, cudaSurfaceObject_t surf, const TgGrid& g, int nxp, int nwx,
                          cudaStream_t st) {
  const size_t tile_bytes = (size_t)(TGE_SEG + 2 * WB) * 32 * sizeof(uint32_t);
  k_edt_xy<WB><<<dim3(tg_div_up(nxp, 64), tg_div_up(g.ny, TGE_SEG), g.nz), TGE_THREADS, tile_bytes, st>>>(occ_bits, g2, g.nx, nxp, nwx, g.ny,
                                                                                                         g.window);
  k_edt_z<WB><<<dim3(tg_div_up(g.nx, 64), g.ny, tg_div_up(g.nz, TGE_SEG)), TGE_THREADS, tile_bytes, st>>>(g2, df, surf, g.nx, nxp, g.ny,
                                                                                                        g.nz, g.window);
}

// occupancy bits of the half-resolution grid from the fine codes (code 0 <=> occupied voxel): one thread per block of 2^3
// voxels, one ballot per 32 blocks.  grid (mnwx, ceil(mny / 8), mnz), block (32, 8).
__global__ void __launch_bounds__(256)
k_mid_bits(const uint8_t* __restrict__ df, int nx, int ny, int nz, uint32_t* __restrict__ bits, int mnx, int mny, int mnwx) {
  const int mx = blockIdx.x * 32 + threadIdx.x, my = blockIdx.y * 8 + threadIdx.y, mz = blockIdx.z;
  if (my >= mny) return;  // warp-uniform (a warp is one row of threadIdx.y)
  bool occ = false;
  if (mx < mnx) {
#pragma unroll
    for (int dz = 0; dz < 2; ++dz)
#pragma unroll
      for (int dy = 0; dy < 2; ++dy) {
        const int z = 2 * mz + dz, y = 2 * my + dy;
        if (z < nz && y < ny) {
          const uint8_t* row = df + ((size_t)z * ny + y) * nx;
          occ |= __ldg(row + 2 * mx) == 0;
          if (2 * mx + 1 < nx) occ |= __ldg(row + 2 * mx + 1) == 0;
        }
      }
  }
  const unsigned w = __ballot_sync(0xffffffffu, occ);
  if (threadIdx.x == 0) bits[((size_t)mz * mny + my) * mnwx + blockIdx.x] = w;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
#ifndef TG_WINDOW_VOX
#define TG_WINDOW_VOX 0
#endif
static const double kEpsFp = 2e-5;  // metres: fp32 slop of centres / voxel index rounding

static void tg_compute_thresholds(const TgGrid& g, TgThr* thr) {
  const double q = g.q, slack = g.slack;
  for (int j = 0; j < TG_N_SPHERES; ++j) {
    const double R = (double)TGM_H_RADIUS_K[j] + (double)g.r_pt;
    int fg = (int)floor((R + slack) / q) + 1;
    int cl = (R - slack) > 0 ? (int)floor((R - slack) / q) - 1 : -1;
    if (cl > 254) cl = 254;
    if (cl < -1) cl = -1;
    thr->free_ge[j] = fg;
    thr->coll_le[j] = cl;
  }
  for (int b = 0; b < 16; ++b) thr->bound_ge[b] = 256;  // never
  for (int b = 0; b < TGM_N_BOUNDS; ++b) {
    const double R = (double)TGM_H_BOUND_RADIUS[b] + (double)g.r_pt;
    thr->bound_ge[b] = (int)floor((R + slack) / q) + 1;  // > 255: the window is too small to ever prove it
  }
  for (int j = 0; j < TG_N_SPHERES; ++j)
    thr->cfree_ge[j] = (int)floor(((double)TGM_H_RADIUS_K[j] + (double)g.r_pt + kEpsFp) / (double)g.cg_q) + 1;
  for (int b = 0; b < 16; ++b) thr->cbound_ge[b] = 256;
  for (int b = 0; b < TGM_N_BOUNDS; ++b)
    thr->cbound_ge[b] = (int)floor(((double)TGM_H_BOUND_RADIUS[b] + (double)g.r_pt + kEpsFp) / (double)g.cg_q) + 1;
  // code c at the source point => no cloud point within c*q - slack of it; the sphere is inside the ball of radius
  // sep + r around the source (sep = centre separation in the rigid frame, rounded up in the generator)
  for (int j = 0; j < TG_N_SPHERES; ++j) {
    const double R = (double)TGM_H_K_SRCSEP[j] + (double)TGM_H_RADIUS_K[j] + (double)g.r_pt;
    thr->src_ge[j] = tgm_k_src(j) < 0 ? 256 : (int)floor((R + slack) / q) + 1;
  }
}

// choose grid geometry for the cloud box [lo, hi]; returns false if it cannot fit
static bool tg_plan_grids(tg_env* env, const float lo[3], const float hi[3], TgGrid* g) {
  double vpm = env->vpm_requested, cpm = env->cpm_requested, cpmx = env->cpmx_requested;
  const double r_need = 0.24 + (double)env->r_pt;  // largest robot sphere + point radius
  for (int it = 0; it < 64; ++it) {
    const double slack = sqrt(3.0) / vpm + kEpsFp;
    int W = (int)ceil((r_need + slack + 0.004) * vpm) + 1;
    if (W < TG_WINDOW_VOX) W = TG_WINDOW_VOX;  // wider than the largest sphere needs: lets the enclosing-sphere probes prove more
    if (W > 31) {  // window too wide for the transform: the grid is finer than useful
      vpm *= 0.9;
      continue;
    }
    int m[3], n[3];
    double nvox = 1;
    bool ok = true;
    for (int a = 0; a < 3; ++a) {
      m[a] = (int)floor((double)lo[a] * vpm) - (W + 3);
      n[a] = (int)ceil((double)hi[a] * vpm) + (W + 3) - m[a] + 1;
      if (n[a] > 1024 || abs(m[a]) > 4000000) ok = false;
      nvox *= n[a];
    }
    if (!ok || nvox > (double)env->max_voxels) {
      vpm *= 0.9;
      if (vpm < 1.0) return false;
      continue;
    }
    g->vpm = (float)vpm;
    vpm = g->vpm;  // what the kernels will really use
    g->window = W;
    g->q = (float)((W / vpm) / 255.0);
    g->slack = (float)(sqrt(3.0) / vpm + kEpsFp + 0.02 * g->q);
    for (int a = 0; a < 3; ++a) {
      g->bias[a] = TG_MAGIC - (float)m[a];
      g->hi[a] = TG_MAGIC + (float)(n[a] - 1);
    }
    g->nx = n[0]; g->ny = n[1]; g->nz = n[2];
    g->nxy = n[0] * n[1];
    g->idx_bias = TG_MAGIC_BITS * (1u + (unsigned)g->nx + (unsigned)g->nxy);
    // coarse far-field grid: the smallest block edge (in voxels) whose grid fits TG_COARSE_MAX bytes
    for (int r = 2; r <= 512; ++r) {
      const long long cn = (long long)tg_div_up(n[0], r) * tg_div_up(n[1], r) * tg_div_up(n[2], r);
      if (cn <= TG_COARSE_MAX) {
        g->cg_r = r;
        g->cg_inv_r = 1.0f / (float)r;
        g->gnx = tg_div_up(n[0], r); g->gny = tg_div_up(n[1], r); g->gnz = tg_div_up(n[2], r);
        break;
      }
    }
    g->cg_q = 0.005f;
    break;
  }
  if (g->nx == 0) return false;
  for (int it = 0; it < 64; ++it) {
    int cm[3], cn[3];
    double nc = 1;
    for (int a = 0; a < 3; ++a) {
      const double ca = a == 0 ? cpmx : cpm;
      cm[a] = (int)floor((double)lo[a] * ca) - 1;
      cn[a] = (int)floor((double)hi[a] * ca) - cm[a] + 2;
      nc *= cn[a];
    }
    if (nc > (double)env->max_cells) {
      cpm *= 0.9;
      cpmx *= 0.9;
      continue;
    }
    g->cpm = (float)cpm;
    g->cpmx = (float)cpmx;
    for (int a = 0; a < 3; ++a) g->cm[a] = cm[a];
    g->cnx = cn[0]; g->cny = cn[1]; g->cnz = cn[2];
    return true;
  }
  return false;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static void tg_layout(tg_env* env, TgGrid* g, TgHeader* h) {
  const size_t nvox = (size_t)g->nx * g->ny * g->nz;
  const size_t ncell = (size_t)g->cnx * g->cny * g->cnz;
  h->off_df = TG_HEADER_BYTES;
  h->off_cells = align256(h->off_df + nvox);
  h->off_pts = align256(h->off_cells + 4 * (ncell + 1));
  h->off_cg = align256(h->off_pts + 16 * (size_t)g->n_points);
  h->bytes = align256(h->off_cg + (size_t)g->gnx * g->gny * g->gnz);
  g->cg = env->blob + h->off_cg;
  g->df = env->blob + h->off_df;
  g->cell_start = (const uint32_t*)(env->blob + h->off_cells);
  g->pts = (const float4*)(env->blob + h->off_pts);
}

// Phase 1 reads the distance grid through the texture path: a 3-D cudaArray is stored block-linear, so the spheres of
// one link (a few cm apart in all three directions) tend to share 32-byte sectors, which a row-major grid only
// offers along x; the unit also does the index arithmetic and the clamp.  The array is a copy of the blob's linear
// grid (which stays the thing that is exported / broadcast), refreshed after every build or import.
static int tg_env_ensure_array(tg_env* env, TgGrid& g) {
  g.df_tex = 0;
  if (g.empty) return TG_OK;
  if (env->df_array == nullptr || env->arr_n[0] != g.nx || env->arr_n[1] != g.ny || env->arr_n[2] != g.nz) {
    if (env->df_tex) cudaDestroyTextureObject(env->df_tex);
    if (env->df_surf) cudaDestroySurfaceObject(env->df_surf);
    if (env->df_array) cudaFreeArray(env->df_array);
    env->df_tex = 0;
    env->df_surf = 0;
    env->df_array = nullptr;
    cudaChannelFormatDesc cd = cudaCreateChannelDesc<unsigned char>();
    TG_CUDA(cudaMalloc3DArray(&env->df_array, &cd, make_cudaExtent((size_t)g.nx, (size_t)g.ny, (size_t)g.nz), cudaArraySurfaceLoadStore));
    cudaResourceDesc rd;
    memset(&rd, 0, sizeof(rd));
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = env->df_array;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    TG_CUDA(cudaCreateTextureObject(&env->df_tex, &rd, &td, nullptr));
    TG_CUDA(cudaCreateSurfaceObject(&env->df_surf, &rd));
    env->arr_n[0] = g.nx; env->arr_n[1] = g.ny; env->arr_n[2] = g.nz;
  }
  g.df_tex = (unsigned long long)env->df_tex;
  for (int a = 0; a < 3; ++a) g.tex_bias[a] = 0.5f - (TG_MAGIC - g.bias[a]);  // bias = 2^23 - m
  return TG_OK;
}

// linear grid -> texture array with surface stores (import path).  One thread = 4 voxels of a row.
__global__ void __launch_bounds__(256) k_grid_to_array(const uint8_t* __restrict__ df, cudaSurfaceObject_t surf, int nx, int ny, int nz) {
  const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y, z = blockIdx.z;
  if (x >= nx) return;
  const uint8_t* src = df + ((size_t)z * ny + y) * nx + x;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (x + k < nx) surf3Dwrite(__ldg(src + k), surf, x + k, y, z);
}

// import path: the array is filled from the blob's linear grid (a build writes it from the Z pass)
static int tg_env_refresh_texture(tg_env* env, cudaStream_t st) {
  TgGrid& g = env->grid;
  const int rc = tg_env_ensure_array(env, g);
  if (rc || g.empty) return rc;
#ifdef TG_IMPORT_SURFACE_KERNEL  // measured: 0.227 ms against 0.152 ms for the 3-D copy below (143 MB blob, tools/adopt_time.py)
  if (env->df_surf) {
    k_grid_to_array<<<dim3(tg_div_up(g.nx, 1024), g.ny, g.nz), 256, 0, st>>>(g.df, env->df_surf, g.nx, g.ny, g.nz);
    TG_CUDA(cudaGetLastError());
    return TG_OK;
  }
#endif
  cudaMemcpy3DParms cp;
  memset(&cp, 0, sizeof(cp));
  cp.srcPtr = make_cudaPitchedPtr((void*)g.df, (size_t)g.nx, (size_t)g.nx, (size_t)g.ny);
  cp.dstArray = env->df_array;
  cp.extent = make_cudaExtent((size_t)g.nx, (size_t)g.ny, (size_t)g.nz);
  cp.kind = cudaMemcpyDeviceToDevice;
  TG_CUDA(cudaMemcpy3DAsync(&cp, st));
  return TG_OK;
}

extern "C" int tg_env_create(tg_env** out, int device, int64_t max_points, float voxels_per_metre, float point_radius,
                             int64_t max_voxels) {
  TG_CHECK_ARG(out != nullptr, "tg_env_create: out is NULL");
  TG_CHECK_ARG(max_points > 0 && max_points < (1ll << 31), "tg_env_create: max_points out of range");
  TG_CHECK_ARG(voxels_per_metre >= 5.f && voxels_per_metre <= 100.f, "tg_env_create: voxels_per_metre must be in [5, 100]");
  TG_CHECK_ARG(point_radius >= 0.f && point_radius < 1.f, "tg_env_create: point_radius out of range");
  int ndev = 0;
  TG_CUDA(cudaGetDeviceCount(&ndev));
  TG_CHECK_ARG(device >= 0 && device < ndev, "tg_env_create: no such CUDA device");
  TG_CUDA(cudaSetDevice(device));
  tg_env* env = new tg_env();
  memset(env, 0, sizeof(*env));
  env->device = device;
  env->timing_events = new std::vector<cudaEvent_t>();
  env->max_points = max_points;
  env->max_voxels = max_voxels > 0 ? max_voxels : (int64_t)64 << 20;
  env->max_cells = env->max_voxels / 4;
  env->vpm_requested = voxels_per_metre;
  // cell rows run along x: 7.1 cm x 7.1 cm rows cut into 1.6 cm cells.  Absolute sizes, not multiples of the voxel: what they
  // trade (rows touched against candidate points read) is set by the robot's sphere radii and the cloud's density
  // (sweeps at 50 and 72 voxels/m: profiles/r01s22_cellshape_sweep*.txt, r01s22_cellshape_c1c2.txt)
  env->cpm_requested = 14.f;
  env->cpmx_requested = 64.f;
  env->r_pt = point_radius;
  env->blob_capacity = TG_HEADER_BYTES + align256(env->max_voxels) + align256(4 * (env->max_cells + 1)) +
                       align256(16 * (size_t)max_points) + align256(TG_COARSE_MAX) + 1024;
  TG_CUDA(cudaMalloc(&env->blob, env->blob_capacity));
  TG_CUDA(cudaMalloc(&env->g1, env->max_voxels));
  TG_CUDA(cudaMalloc(&env->cbits, TG_COARSE_MAX / 8 + 4 * 4096));   // coarse occupancy bits (rows padded to words)
  TG_CUDA(cudaMalloc(&env->g2c, 2 * TG_COARSE_MAX + 2 * 65536));    // u16 [gnz][gny][gnx rounded up to even]
  TG_CUDA(cudaMalloc(&env->g2, 2 * env->max_voxels + env->max_voxels / 8 + 4096));  // u16 [nz][ny][nxp], nxp = nx rounded up to even
  TG_CUDA(cudaMalloc(&env->cell_count, 4 * (env->max_cells + 1)));
  TG_CUDA(cudaMalloc(&env->rank, 4 * max_points));
  TG_CUDA(cudaMalloc(&env->sort_tmp, 16 * (size_t)max_points));  // bucket-partitioned points of the staged sort
  TG_CUDA(cudaMalloc(&env->key, 4 * max_points));
  TG_CUDA(cudaMalloc(&env->scan_tmp, 4 * (env->max_cells / SCAN_ITEMS + 2)));
  TG_CUDA(cudaMalloc(&env->bounds_dev, 8 * sizeof(int)));
  TG_CUDA(cudaMallocHost(&env->bounds_host, 8 * sizeof(int)));
  TG_CUDA(cudaMallocHost(&env->header_host, TG_HEADER_BYTES));
  TG_CUDA(cudaMalloc(&env->extras_dev, sizeof(TgExtras)));
  TG_CUDA(cudaMemset(env->extras_dev, 0, sizeof(TgExtras)));
  env->queue_capacity = 1u << 23;  // 8 Mi items = 512 MB (two halves for the host pipeline, one piece for the device API)
  TG_CUDA(cudaMalloc(&env->items, (size_t)env->queue_capacity * sizeof(TgItem)));
  TG_CUDA(cudaMalloc(&env->queue_count, 64));
  TG_CUDA(cudaMalloc(&env->counters_dev, sizeof(TgCounters)));
  TG_CUDA(cudaMemset(env->counters_dev, 0, sizeof(TgCounters)));
  TG_CUDA(cudaMallocHost(&env->counters_host, sizeof(TgCounters)));
  cudaDeviceProp prop;
  TG_CUDA(cudaGetDeviceProperties(&prop, device));
  env->sm_count = prop.multiProcessorCount;
  {
    const int rc = tg_query_occupancy(env);
    if (rc) return rc;
  }
  // empty world until the first build
  env->grid.empty = 1;
  env->grid.r_pt = point_radius;
  *out = env;
  return TG_OK;
}

extern "C" int tg_env_destroy(tg_env* env) {
  if (!env) return TG_OK;
  cudaSetDevice(env->device);
  cudaFree(env->cbits); cudaFree(env->g2c); cudaFree(env->mid);
  cudaFree(env->blob); cudaFree(env->g1); cudaFree(env->g2); cudaFree(env->cell_count);
  cudaFree(env->sort_tmp);
  cudaFree(env->rank); cudaFree(env->key); cudaFree(env->scan_tmp); cudaFree(env->bounds_dev);
  cudaFree(env->xyz_stage); cudaFree(env->extras_dev); cudaFree(env->items); cudaFree(env->queue_count);
  cudaFree(env->counters_dev);
  cudaFreeHost(env->bounds_host); cudaFreeHost(env->header_host); cudaFreeHost(env->counters_host);
  for (int i = 0; i < 2; ++i) {
    if (env->streams[i]) cudaStreamDestroy(env->streams[i]);
    if (env->events[i]) cudaEventDestroy(env->events[i]);
    cudaFreeHost(env->pin_in[i]); cudaFreeHost(env->pin_out[i]); cudaFreeHost(env->pin_clear[i]);
    cudaFree(env->dev_in[i]); cudaFree(env->dev_out[i]); cudaFree(env->dev_clear[i]);
  }
  if (env->df_tex) cudaDestroyTextureObject(env->df_tex);
  if (env->df_surf) cudaDestroySurfaceObject(env->df_surf);
  if (env->df_array) cudaFreeArray(env->df_array);
  cudaFree(env->dev_checks);
  cudaFreeHost(env->pin_checks);
  cudaFreeHost(env->pin_small);
  cudaFreeHost(env->pin_small_out);
  if (env->timing_events) {
    for (cudaEvent_t e : *env->timing_events) cudaEventDestroy(e);
    delete env->timing_events;
  }
  delete env;
  return TG_OK;
}

extern "C" int tg_env_set_bounds(tg_env* env, const float* lo, const float* hi) {
  TG_CHECK_ARG(env != nullptr, "tg_env_set_bounds: env is NULL");
  if (!lo || !hi) {
    env->fixed_bounds = false;
    return TG_OK;
  }
  for (int a = 0; a < 3; ++a) {
    TG_CHECK_ARG(isfinite(lo[a]) && isfinite(hi[a]) && hi[a] > lo[a], "tg_env_set_bounds: need finite lo < hi");
    env->lo[a] = lo[a];
    env->hi[a] = hi[a];
  }
  env->fixed_bounds = true;
  return TG_OK;
}

extern "C" int tg_env_set_cells_per_metre(tg_env* env, float cpm) {
  TG_CHECK_ARG(env != nullptr && cpm >= 2.f && cpm <= 100.f, "tg_env_set_cells_per_metre: need 2 <= cells_per_metre <= 100");
  env->cpm_requested = cpm;
  env->cpmx_requested = cpm;
  return TG_OK;
}

extern "C" int tg_env_set_cell_shape(tg_env* env, float cpm_x, float cpm_yz) {
  TG_CHECK_ARG(env != nullptr && cpm_x >= 2.f && cpm_x <= 200.f && cpm_yz >= 2.f && cpm_yz <= 100.f,
               "tg_env_set_cell_shape: need 2 <= cells_per_metre_x <= 200 and 2 <= cells_per_metre_yz <= 100");
  env->cpmx_requested = cpm_x;
  env->cpm_requested = cpm_yz;
  return TG_OK;
}

extern "C" int tg_env_set_point_radius(tg_env* env, float r) {
  TG_CHECK_ARG(env != nullptr && r >= 0.f && r < 1.f, "tg_env_set_point_radius: bad argument");
  env->r_pt = r;
  return TG_OK;
}

extern "C" int tg_env_build(tg_env* env, const float* xyz, int64_t n, void* stream_) {
  TG_CHECK_ARG(env != nullptr, "tg_env_build: env is NULL");
  TG_CHECK_ARG(n >= 0 && (n == 0 || xyz != nullptr), "tg_env_build: bad cloud");
  if (n > env->max_points) {
    tg_set_error("tg_env_build: %lld points exceed the env capacity %lld", (long long)n, (long long)env->max_points);
    return TG_E_CAPACITY;
  }
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(env->device));
  TgGrid g;
  memset(&g, 0, sizeof(g));
  g.r_pt = env->r_pt;
  float lo[3], hi[3];
  bool have_box = false;
  if (n > 0) {
    if (env->fixed_bounds) {
      memcpy(lo, env->lo, sizeof(lo));
      memcpy(hi, env->hi, sizeof(hi));
      have_box = true;
    } else {
      k_bounds_init<<<1, 32, 0, st>>>(env->bounds_dev);
      k_bounds<<<min(tg_div_up(n, 256), 4 * env->sm_count), 256, 0, st>>>(xyz, (int)n, env->bounds_dev);
      TG_CUDA(cudaMemcpyAsync(env->bounds_host, env->bounds_dev, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
      TG_CUDA(cudaStreamSynchronize(st));
      env->launches += 2;
      if (env->bounds_host[0] <= env->bounds_host[3]) {  // at least one finite point
        for (int a = 0; a < 3; ++a) {
          lo[a] = ord2f(env->bounds_host[a]);
          hi[a] = ord2f(env->bounds_host[3 + a]);
        }
        have_box = true;
      }
    }
  }
  if (!have_box) {  // empty world: every configuration is free of the cloud
    g.empty = 1;
    env->grid = g;
    memset(&env->thr, 0, sizeof(env->thr));
    TgHeader* h = env->header_host;
    memset(h, 0, sizeof(TgHeader));
    h->magic = 0x54594752495443ull;
    h->bytes = TG_HEADER_BYTES;
    h->grid = g;
    TG_CUDA(cudaMemcpyAsync(env->blob, h, sizeof(TgHeader), cudaMemcpyHostToDevice, st));
    return TG_OK;
  }
  if (!tg_plan_grids(env, lo, hi, &g)) {
    tg_set_error("tg_env_build: cloud box [%g,%g,%g]-[%g,%g,%g] does not fit the grid capacity", lo[0], lo[1], lo[2],
                 hi[0], hi[1], hi[2]);
    return TG_E_CAPACITY;
  }
  g.n_points = (int)n;
  TgHeader* h = env->header_host;
  memset(h, 0, sizeof(TgHeader));
  h->magic = 0x54594752495443ull;
  tg_layout(env, &g, h);
  if (h->bytes > env->blob_capacity) {
    tg_set_error("tg_env_build: structure needs %zu bytes, capacity %zu", (size_t)h->bytes, env->blob_capacity);
    return TG_E_CAPACITY;
  }
  const size_t nvox = (size_t)g.nx * g.ny * g.nz;
  const int ncell = g.cnx * g.cny * g.cnz;
  uint8_t* df = env->blob + h->off_df;
  uint32_t* cell_start = (uint32_t*)(env->blob + h->off_cells);
  float4* pts = (float4*)(env->blob + h->off_pts);

  {
    const int rc = tg_env_ensure_array(env, g);  // (re)allocated only when the grid dims change
    if (rc) return rc;
  }
  const cudaSurfaceObject_t surf = env->df_surf;
  const int nwx = (g.nx + 31) / 32, nxp = (g.nx + 1) & ~1;
  uint32_t* occ_bits = (uint32_t*)env->g1;  // occupancy bitmask [nz][ny][nwx]
  TG_CUDA(cudaMemsetAsync(occ_bits, 0, 4 * (size_t)g.nz * g.ny * nwx, st));
  const int cnwx = (g.gnx + 31) / 32;
  TG_CUDA(cudaMemsetAsync(env->cbits, 0, 4 * (size_t)g.gnz * g.gny * cnwx, st));
#ifdef TG_SORT_GLOBAL_ATOMICS
  TG_CUDA(cudaMemsetAsync(env->cell_count, 0, 4 * ((size_t)ncell + 1), st));
#else
  TG_CUDA(cudaMemsetAsync(env->cell_count, 0, 4 * (size_t)tg_div_up(ncell, 2048), st));  // bucket counts only (>= the bucket count)
#endif
  TG_CUDA(cudaMemsetAsync(env->bounds_dev + 6, 0, sizeof(int), st));
#ifdef TG_SORT_GLOBAL_ATOMICS
  k_count<<<tg_div_up(n, 256), 256, 0, st>>>(xyz, (int)n, g, occ_bits, nwx, env->cbits, cnwx, env->cell_count, env->key, env->rank,
                                             env->bounds_dev + 6);
  const int nb = tg_div_up(ncell, SCAN_ITEMS);
  k_scan_block_sums<<<nb, 256, 0, st>>>(env->cell_count, ncell, env->scan_tmp);
  k_scan_sums<<<1, 1024, 0, st>>>(env->scan_tmp, nb);
  k_scan_apply<<<nb, 256, 0, st>>>(env->cell_count, ncell, env->scan_tmp, cell_start);
  k_scatter<<<tg_div_up(n, 256), 256, 0, st>>>(xyz, (int)n, env->key, env->rank, cell_start, pts);
#else
  {
    int cpb_shift = 11;  // 2048 cells per bucket: many small blocks in pass D; doubled for very large cell grids
    while ((1 << cpb_shift) < TGS_CPB_MAX && tg_div_up(ncell, 1 << cpb_shift) > TGS_MAX_BUCKETS) ++cpb_shift;
    const int nbuckets = tg_div_up(ncell, 1 << cpb_shift);
    if (nbuckets > TGS_MAX_BUCKETS) {
      tg_set_error("tg_env_build: %d cells need more sort buckets than supported (coarsen the cells)", ncell);
      return TG_E_CAPACITY;
    }
    uint32_t* bucket_count = env->cell_count;               // [nbuckets] counts, then [nbuckets + 1] starts, then [nbuckets] cursors
    uint32_t* bucket_start = bucket_count + nbuckets;
    uint32_t* bucket_cursor = bucket_start + nbuckets + 1;
    float4* tmp = reinterpret_cast<float4*>(env->sort_tmp);
    const int tiles = tg_div_up(n, TGS_TILE);
    k_sort_hist<<<tiles, 256, 4 * nbuckets, st>>>(xyz, (int)n, g, occ_bits, nwx, env->cbits, cnwx, nbuckets, cpb_shift, env->key,
                                                   bucket_count, env->bounds_dev + 6);
    k_sort_scan<<<1, 1024, 0, st>>>(bucket_count, nbuckets, bucket_start, bucket_cursor);
    k_sort_partition<<<tiles, 256, 8 * nbuckets, st>>>(xyz, (int)n, env->key, nbuckets, cpb_shift, bucket_cursor, tmp);
    k_sort_bucket<<<nbuckets, 256, 4 << cpb_shift, st>>>(tmp, bucket_start, nbuckets, cpb_shift, ncell, cell_start, pts);
  }
#endif
  switch ((g.window + 3) / 4) {  // window rounded up to a multiple of 4 (W <= 31)
    case 0: case 1: case 2: tg_launch_edt<8>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    case 3: tg_launch_edt<12>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    case 4: tg_launch_edt<16>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    case 5: tg_launch_edt<20>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    case 6: tg_launch_edt<24>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    case 7: tg_launch_edt<28>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
    default: tg_launch_edt<32>(occ_bits, (uint32_t*)env->g2, df, surf, g, nxp, nwx, st); break;
  }
  {  // coarse far-field grid: lower bounds out to ~1.2 m (255 codes of 5 mm)
    const float cell = (float)g.cg_r / g.vpm;
    const int Wc = min(31, (int)ceilf((255.0f * g.cg_q + 1.7320509f * cell) / cell) + 1);
    uint8_t* cg = env->blob + h->off_cg;
    switch ((Wc + 3) / 4) {
      case 0: case 1: case 2: tg_launch_edt_coarse<8>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      case 3: tg_launch_edt_coarse<12>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      case 4: tg_launch_edt_coarse<16>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      case 5: tg_launch_edt_coarse<20>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      case 6: tg_launch_edt_coarse<24>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      case 7: tg_launch_edt_coarse<28>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
      default: tg_launch_edt_coarse<32>(env->cbits, env->g2c, cg, g, Wc, cell, st); break;
    }
  }
  TG_CUDA(cudaGetLastError());
  env->launches += 9;

  tg_compute_thresholds(g, &env->thr);
  env->grid = g;
  env->mid_valid = 0;
  h->grid = g;
  h->thr = env->thr;
  TG_CUDA(cudaMemcpyAsync(env->blob, h, sizeof(TgHeader), cudaMemcpyHostToDevice, st));
  return TG_OK;
}

// The half-resolution grid of the clearance query, (re)built on `st` when the structure changed since the last call.
// Scratch: g1 (occupancy bits) and g2 (pass-Y output) of the build.
int tg_env_ensure_mid(tg_env* env, cudaStream_t st) {
  if (env->mid_valid || env->grid.empty) return TG_OK;
  const TgGrid& g = env->grid;
  TgGrid gm = g;
  gm.nx = (g.nx + 1) / 2; gm.ny = (g.ny + 1) / 2; gm.nz = (g.nz + 1) / 2;
  gm.window = 31;
  const size_t need = (size_t)gm.nx * gm.ny * gm.nz;
  if (need > env->mid_capacity) {
    if (env->mid) TG_CUDA(cudaFree(env->mid));
    env->mid = nullptr;
    env->mid_capacity = 0;
    TG_CUDA(cudaMalloc(&env->mid, need + need / 8 + 4096));
    env->mid_capacity = need + need / 8 + 4096;
  }
  const int nwx = (gm.nx + 31) / 32, nxp = (gm.nx + 1) & ~1;
  uint32_t* bits = (uint32_t*)env->g1;
  k_mid_bits<<<dim3(nwx, tg_div_up(gm.ny, 8), gm.nz), dim3(32, 8), 0, st>>>(g.df, g.nx, g.ny, g.nz, bits, gm.nx, gm.ny, nwx);
  tg_launch_edt<32>(bits, (uint32_t*)env->g2, env->mid, 0, gm, nxp, nwx, st);
  TG_CUDA(cudaGetLastError());
  // once per structure: the grid must be complete before ANY stream of this handle reads it (the host pipeline alternates
  // between two streams) and before a later build reuses the scratch buffers
  TG_CUDA(cudaStreamSynchronize(st));
  env->launches += 3;
  env->midg.mg = env->mid;
  env->midg.nx = gm.nx; env->midg.ny = gm.ny; env->midg.nz = gm.nz;
  env->midg.q = (float)((31.0 * 2.0 / (double)g.vpm) / 255.0);
  env->midg.slack = (float)(2.0 * sqrt(3.0) / (double)g.vpm + kEpsFp + 0.02 * env->midg.q);
  env->mid_valid = 1;
  return TG_OK;
}

extern "C" int tg_env_build_host(tg_env* env, const float* xyz, int64_t n, void* stream_) {
  TG_CHECK_ARG(env != nullptr, "tg_env_build_host: env is NULL");
  TG_CHECK_ARG(n >= 0 && (n == 0 || xyz != nullptr), "tg_env_build_host: bad cloud");
  if (n > env->max_points) {
    tg_set_error("tg_env_build_host: %lld points exceed the env capacity %lld", (long long)n, (long long)env->max_points);
    return TG_E_CAPACITY;
  }
  TG_CUDA(cudaSetDevice(env->device));
  cudaStream_t st = (cudaStream_t)stream_;
  if (n > 0) {
    if (!env->xyz_stage) TG_CUDA(cudaMalloc(&env->xyz_stage, 12 * (size_t)env->max_points));
    TG_CUDA(cudaMemcpyAsync(env->xyz_stage, xyz, 12 * (size_t)n, cudaMemcpyHostToDevice, st));
  }
  return tg_env_build(env, env->xyz_stage, n, stream_);
}

extern "C" int tg_env_stats(const tg_env* env, double* out8) {
  TG_CHECK_ARG(env != nullptr && out8 != nullptr, "tg_env_stats: bad argument");
  const TgGrid& g = env->grid;
  out8[0] = g.n_points;
  out8[1] = 0;  // filled by tg_env_counters-style readback below if available
  out8[2] = g.nx; out8[3] = g.ny; out8[4] = g.nz;
  out8[5] = (double)g.cnx * g.cny * g.cnz;
  out8[6] = g.vpm;
  out8[7] = g.slack;
  if (!g.empty) {
    int dropped = 0;
    if (cudaMemcpy(&dropped, env->bounds_dev + 6, sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess) out8[1] = dropped;
  }
  return TG_OK;
}

// Layout of the built structure inside the blob (for tests and tools that read the grids back).
extern "C" int tg_env_grid_info(const tg_env* env, double* out20) {
  TG_CHECK_ARG(env != nullptr && out20 != nullptr, "tg_env_grid_info: bad argument");
  const TgGrid& g = env->grid;
  const TgHeader* h = env->header_host;
  for (int i = 0; i < 20; ++i) out20[i] = 0;
  out20[0] = g.empty;
  if (g.empty) return TG_OK;
  out20[1] = g.nx; out20[2] = g.ny; out20[3] = g.nz;
  for (int a = 0; a < 3; ++a) out20[4 + a] = g.bias[a];
  out20[7] = g.vpm; out20[8] = g.q; out20[9] = g.slack; out20[10] = g.window;
  out20[11] = (double)h->off_df; out20[12] = (double)h->off_cg;
  out20[13] = g.gnx; out20[14] = g.gny; out20[15] = g.gnz; out20[16] = g.cg_r; out20[17] = g.cg_q;
  return TG_OK;
}

extern "C" size_t tg_env_blob_capacity(const tg_env* env) { return env ? env->blob_capacity : 0; }

extern "C" int tg_env_blob_ptr(tg_env* env, void** blob) {
  TG_CHECK_ARG(env != nullptr && blob != nullptr, "tg_env_blob_ptr: bad argument");
  *blob = env->blob;
  return TG_OK;
}

extern "C" int tg_env_export(tg_env* env, void** blob, size_t* bytes) {
  TG_CHECK_ARG(env != nullptr && blob != nullptr && bytes != nullptr, "tg_env_export: bad argument");
  *blob = env->blob;
  *bytes = (size_t)env->header_host->bytes;
  if (*bytes == 0) {
    tg_set_error("tg_env_export: nothing built yet");
    return TG_E_STATE;
  }
  return TG_OK;
}

extern "C" int tg_env_import(tg_env* env, const void* blob, size_t bytes, void* stream_) {
  TG_CHECK_ARG(env != nullptr && blob != nullptr && bytes >= TG_HEADER_BYTES, "tg_env_import: bad argument");
  if (bytes > env->blob_capacity) {
    tg_set_error("tg_env_import: blob of %zu bytes exceeds capacity %zu", bytes, env->blob_capacity);
    return TG_E_CAPACITY;
  }
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(env->device));
  if (blob != env->blob) TG_CUDA(cudaMemcpyAsync(env->blob, blob, bytes, cudaMemcpyDeviceToDevice, st));
  TG_CUDA(cudaMemcpyAsync(env->header_host, env->blob, sizeof(TgHeader), cudaMemcpyDeviceToHost, st));
  TG_CUDA(cudaStreamSynchronize(st));
  TgHeader* h = env->header_host;
  if (h->magic != 0x54594752495443ull || h->bytes > bytes) {
    tg_set_error("tg_env_import: not a tygrit_b200 structure blob");
    return TG_E_ARG;
  }
  TgGrid g = h->grid;
  if (!g.empty) {
    g.df = env->blob + h->off_df;
    g.cell_start = (const uint32_t*)(env->blob + h->off_cells);
    g.pts = (const float4*)(env->blob + h->off_pts);
    g.cg = env->blob + h->off_cg;
  }
  env->grid = g;
  env->mid_valid = 0;
  env->thr = h->thr;
  env->r_pt = g.r_pt;
  return tg_env_refresh_texture(env, st);
}

extern "C" int tg_env_set_obstacles(tg_env* env, const float* xyzr, int n, void* stream_) {
  TG_CHECK_ARG(env != nullptr && n >= 0 && n <= TG_MAX_OBSTACLES && (n == 0 || xyzr), "tg_env_set_obstacles: bad argument (max 64 spheres)");
  env->extras.n_obs = n;
  for (int i = 0; i < n; ++i) env->extras.obs[i] = make_float4(xyzr[4 * i], xyzr[4 * i + 1], xyzr[4 * i + 2], xyzr[4 * i + 3]);
  TG_CUDA(cudaSetDevice(env->device));
  TG_CUDA(cudaMemcpyAsync(env->extras_dev, &env->extras, sizeof(TgExtras), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
  TG_CUDA(cudaStreamSynchronize((cudaStream_t)stream_));
  return TG_OK;
}

extern "C" int tg_env_set_attachment(tg_env* env, const float* xyzr, int n, const float* pos, const float* quat, void* stream_) {
  TG_CHECK_ARG(env != nullptr && n >= 0 && n <= TG_MAX_ATTACHED && (n == 0 || xyzr), "tg_env_set_attachment: bad argument (max 64 spheres)");
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 0};
  if (pos) for (int a = 0; a < 3; ++a) t[a] = pos[a];
  if (quat) {
    double x = quat[0], y = quat[1], z = quat[2], w = quat[3];
    const double nn = sqrt(x * x + y * y + z * z + w * w);
    TG_CHECK_ARG(nn > 0, "tg_env_set_attachment: zero quaternion");
    x /= nn; y /= nn; z /= nn; w /= nn;
    R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w); R[2] = 2 * (x * z + y * w);
    R[3] = 2 * (x * y + z * w); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
    R[6] = 2 * (x * z - y * w); R[7] = 2 * (y * z + x * w); R[8] = 1 - 2 * (x * x + y * y);
  }
  env->extras.n_att = n;
  for (int i = 0; i < n; ++i) {
    const double px = xyzr[4 * i], py = xyzr[4 * i + 1], pz = xyzr[4 * i + 2];
    env->extras.att[i] = make_float4((float)(R[0] * px + R[1] * py + R[2] * pz + t[0]), (float)(R[3] * px + R[4] * py + R[5] * pz + t[1]),
                                     (float)(R[6] * px + R[7] * py + R[8] * pz + t[2]), xyzr[4 * i + 3]);
  }
  TG_CUDA(cudaSetDevice(env->device));
  TG_CUDA(cudaMemcpyAsync(env->extras_dev, &env->extras, sizeof(TgExtras), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
  TG_CUDA(cudaStreamSynchronize((cudaStream_t)stream_));
  return TG_OK;
}
