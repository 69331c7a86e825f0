// common.cuh — shared types of the tygrit_b200 CUDA library (sm_100a).
//
// Data layout in HBM (one collision world = one tg_env):
//   blob = [ TgHeader (4 KB) | df codes u8 [nz][ny][nx] | cell_start u32 [ncells+1] | pts float4 [n] ]
// * df: "distance-bound grid".  Voxel i along an axis is centred on (i + m) / vpm metres.
//   code k < 255 means the distance from the voxel CENTRE to the nearest OCCUPIED voxel centre is
//   in [k q, (k+1) q); k = 255 means >= 255 q = W / vpm.  With slack = sqrt(3)/vpm + eps this
//   bounds dist(x, cloud) for every x in the voxel:  k q - slack <= dist <= (k+1) q + slack.
// * cell grid: points counting-sorted by cell key ((iz*cny + iy)*cnx + ix), x fastest, so one row
//   of cells is one contiguous point range — the exact narrow phase reads it linearly.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/tygrit_b200.h"

#define TG_MAGIC 8388608.0f        // 2^23: float -> int by mantissa trick
#define TG_MAGIC_BITS 0x4B000000u  // bit pattern of 2^23
#define TG_HEADER_BYTES 4096
#define TG_CLEARANCE_CAP 0.25f
#define TG_FLAG_SELF 1  // tg_env.flags: self-collision term on (tg_env_set_self_collision)
#define TG_HOST_MAX_CHUNKS 4096
#ifndef TG_COARSE_MAX
#define TG_COARSE_MAX 114688  // bytes of the coarse far-field grid: it lives in SHARED memory of the phase-1 kernel next to the 80 KB
                              // ring; 112 + 80 KB keep the SM's shared-memory carve-out at 196 KB (228 KB would leave 28 KB of L1: phase 1 + 11 %)
#endif
#define TG_HOST_SMALL 16384  // host batches up to this size go through one packed H2D copy and one D2H copy

struct TgGrid {
  // distance-bound grid
  float vpm;            // voxels per metre
  float bias[3];        // 2^23 - m  (exact integers as floats)
  float hi[3];          // 2^23 + n - 1
  int nx, ny, nz;
  int nxy;              // nx * ny
  unsigned idx_bias;    // TG_MAGIC_BITS * (1 + nx + nxy)  (mod 2^32)
  float q;              // metres per code unit
  float slack;          // metres
  int window;           // W (voxels)
  // narrow-phase cell grid
  float cpm;            // cells per metre along y and z
  float cpmx;           // cells per metre along x (the row direction: finer x cells tighten the chord, coarser y/z cells mean fewer rows)
  int cm[3];            // cell index = floor(p * cpm{x}) - cm
  int cnx, cny, cnz;
  float r_pt;           // cloud point radius
  int n_points;
  int n_dropped;
  int empty;            // 1 = no cloud geometry
  // device pointers (into the blob)
  const uint8_t* df;
  unsigned long long df_tex;  // cudaTextureObject_t over a 3-D array copy of df (process-local, set after build / import)
  float tex_bias[3];          // voxel coordinate -> texel coordinate: 0.5 - m
  const uint32_t* cell_start;
  const float4* pts;
  // coarse far-field grid (one byte per block of cg_r^3 voxels): code c means dist(x, cloud) >= c * cg_q for every x of the
  // block.  Small enough (<= TG_COARSE_MAX bytes) to be copied into shared memory by every CTA of the phase-1 kernel.
  const uint8_t* cg;
  int gnx, gny, gnz;    // coarse dims = ceil(n / cg_r)
  int cg_r;             // voxels per coarse block edge
  float cg_inv_r;       // 1 / cg_r
  float cg_q;           // metres per coarse code unit
};

// per-sphere integer thresholds on the df code (kernel order): definitely free iff code >= free_ge,
// definitely colliding iff code <= coll_le (coll_le = -1: never).
struct TgThr {
  int free_ge[TG_N_SPHERES];
  int coll_le[TG_N_SPHERES];
  int bound_ge[16];  // per cull group (fetch_model.cuh TGM_BOUND): enclosing sphere definitely free iff code >= bound_ge
  int src_ge[TG_N_SPHERES];  // derived bound: sphere definitely free iff the code at its source point (tgm_k_src) >= src_ge
  int cfree_ge[TG_N_SPHERES];  // coarse far-field code at the sphere centre >= cfree_ge: definitely free (no fine probe needed)
  int cbound_ge[16];           // same for the enclosing sphere of a cull group
};

// Half-resolution distance-bound grid of the CLEARANCE query (built lazily from the fine grid by tg_env_ensure_mid, not part
// of the blob): one code per block of 2^3 voxels, window 31 blocks (0.69 m at 90 voxels/m, where the fine grid saturates at
// 0.31 m).  code k < 255: dist(x, cloud) in [k q - slack, (k + 1) q + slack] for every x of the block; 255: >= 255 q - slack.
struct TgMid {
  const uint8_t* mg;  // [nz][ny][nx], block (i, j, k) = voxels 2i..2i+1, ...
  int nx, ny, nz;
  float q, slack;
};

struct TgExtras {  // sphere obstacles + end-effector attachment (rarely used; brute force)
  int n_obs;
  int n_att;
  float4 obs[TG_MAX_OBSTACLES];  // world xyz, r
  float4 att[TG_MAX_ATTACHED];   // gripper_link-frame xyz (offset pose already applied), r
};

struct TgHeader {  // first TG_HEADER_BYTES of the blob
  uint64_t magic;
  uint64_t bytes;  // used bytes of the blob
  uint64_t off_df, off_cells, off_pts, off_cg;
  TgGrid grid;     // pointers are rebased on import
  TgThr thr;
};
static_assert(sizeof(TgHeader) <= TG_HEADER_BYTES, "header too large");

// queue item of the exact narrow phase: one configuration with >= 1 ambiguous sphere
struct __align__(16) TgItem {
  uint32_t target;    // index into the verdict array
  uint32_t mask[4];   // ambiguous spheres (kernel order), bit j
  float cfg[11];      // x, y, theta, q0..q7
};
static_assert(sizeof(TgItem) == 64, "item is one 64-byte line");

struct TgQueue {
  TgItem* items;
  unsigned* count;
  unsigned capacity;
};

struct TgCounters {
  unsigned long long definite_collisions;
  unsigned long long items;
  unsigned long long overflow_inline;
  unsigned long long narrow_points;
};

void tg_set_error(const char* fmt, ...);

#define TG_CUDA(expr)                                                                  \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      tg_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return TG_E_CUDA;                                                                \
    }                                                                                  \
  } while (0)

#define TG_CHECK_ARG(cond, msg)   \
  do {                            \
    if (!(cond)) {                \
      tg_set_error("%s", msg);    \
      return TG_E_ARG;            \
    }                             \
  } while (0)

struct tg_env {
  int device;
  int64_t max_points;
  int64_t max_voxels;
  int64_t max_cells;
  float vpm_requested;
  float cpm_requested;   // y / z
  float cpmx_requested;  // x
  float r_pt;
  int window;
  bool fixed_bounds;
  float lo[3], hi[3];
  // blob + scratch
  uint8_t* blob;
  size_t blob_capacity;
  cudaArray_t df_array;       // block-linear copy of the distance grid, read through the texture path by phase 1
  cudaTextureObject_t df_tex;
  cudaSurfaceObject_t df_surf;  // the Z pass of the build writes the array directly
  int arr_n[3];
  uint32_t* cbits;       // coarse occupancy bitmask [gnz][gny][ceil(gnx/32)]
  uint32_t* g2c;         // coarse EDT intermediate, u16 pairs
  uint8_t* g1;           // EDT pass X output, u8 [nvox]
  uint16_t* g2;          // EDT pass Y output, u16 [nvox]
  uint32_t* cell_count;  // [max_cells + 1]
  uint32_t* rank;        // [max_points]
  void* sort_tmp;        // float4 [max_points]: points grouped by sort bucket (x, y, z, cell key)
  uint32_t* key;         // [max_points]
  uint32_t* scan_tmp;    // block sums
  float* xyz_stage;      // device staging for host clouds [max_points*3]
  int* bounds_dev;       // 6 ordered ints + dropped counter
  int* bounds_host;      // pinned
  TgHeader* header_host; // pinned mirror of the blob header
  // query side
  TgGrid grid;
  TgThr thr;
  TgExtras extras;
  TgExtras* extras_dev;
  TgItem* items;
  unsigned* queue_count;
  unsigned queue_capacity;
  unsigned queue_limit;  // test hook: 0 = use the whole queue
  int flags;             // TG_FLAG_*: optional terms evaluated by their own kernels after the cloud verdicts
  uint8_t* mid;          // half-resolution grid of the clearance query (lazy; TgMid)
  size_t mid_capacity;
  int mid_valid;         // 0 after every build / import
  TgMid midg;
  int no_coarse;         // 1 = never use the coarse far-field variant of the configuration kernel (test / tuning hook)
  TgCounters* counters_dev;
  TgCounters* counters_host;  // pinned
  int64_t launches;
  int timing;                                  // record CUDA events around check-config launches
  std::vector<cudaEvent_t>* timing_events;     // triples: start, after phase 1, after phase 2
  // host-buffer pipeline
  cudaStream_t streams[2];
  cudaEvent_t events[2];
  float* pin_in[2];
  uint8_t* pin_out[2];
  float* dev_in[2];
  uint8_t* dev_out[2];
  float* dev_clear[2];
  float* pin_clear[2];
  int64_t chunk;
  int64_t* dev_checks;  // per-chunk sample counters of the host edge pipeline (TG_HOST_MAX_CHUNKS entries)
  int64_t* pin_checks;
  float* pin_small;      // pinned staging of the small-batch host path (one H2D, one D2H per call)
  uint8_t* pin_small_out;
  int sm_count;
  int occ_configs, occ_resolve, occ_clear;  // resident blocks per SM of the persistent query kernels
};

int tg_check_configs_slot(tg_env* env, const float* arm, const float* base, int64_t B, uint8_t* valid, float* clearance,
                          void* stream, int slot);
int tg_check_edges_slot(tg_env* env, const float* arm0, const float* base0, const float* arm1, const float* base1, int64_t E,
                        int rake, float resolution, uint8_t* valid, int64_t* n_checks, void* stream, int slot);

int tg_query_occupancy(tg_env* env);
int tg_env_ensure_mid(tg_env* env, cudaStream_t st);

static inline int tg_div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
