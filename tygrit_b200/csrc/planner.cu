// planner.cu — device-resident planners on top of tg_check_edges / tg_check_configs.
//
// The reference plans inside third-party VAMP (planning/fetch/algorithms.py:23-232: RRT-Connect for the arm,
// multilayer_rrtc = hybrid-A* base layer + arm RRT-Connect, FCIT* = lazily evaluated fully connected graph), whose
// source is absent; what TyGrit fixes is the surface and the parameters (reverse_penalty = 50, batch_size = 4096,
// max_samples = 8192, XY bounds from the cloud, interpolate(16) / interpolate(0.08)).  These are B200 designs of the
// same planners: every data structure (trees, lattice, roadmap, candidate edges, verdict bytes) lives in HBM, every
// step is a kernel, and the host only enqueues and polls a status word in mapped pinned memory - no host<->device
// copy inside a planning loop.
//
// Configuration vector everywhere: q[11] = [x, y, theta, q0..q7], theta is an angle (distances and steering use the
// shortest arc; a target's theta is stored UNWRAPPED relative to its source so that tg_check_edges, which interpolates
// linearly, follows that arc).
#include <math.h>
#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <chrono>
#include <vector>

#include "common.cuh"

#define TGP_DIM 11
#define TGP_STRIDE 12  // floats per stored node (16-byte rows)
#define TGP_PI 3.14159265358979f
#define TGP_MAX_PATH 256

struct TgPlanStatus {  // mapped pinned memory: written by kernels, polled by the host after an event
  int solved;
  int tree_of_a;       // which tree (0 = start tree, 1 = goal tree) was being extended when the trees met
  int node_a, node_b;  // connection: node in the extended tree, node in the other tree
  int iterations;
  int n_path;
  int changed;         // relaxation sweeps: something improved
  int aux;
};

struct tg_planner {
  tg_env* env;
  int device;
  int cap, batch;
  float* nodes[2];   // [cap][TGP_STRIDE]
  int* parent[2];    // [cap]
  int* count;        // [2] device
  // candidate edges of one iteration (extension, then connection)
  float *arm0, *base0, *arm1, *base1;  // [batch] x 8 / 3
  uint8_t* valid;                      // [batch]
  int* near_idx;                       // [batch]
  int *conn_a, *conn_b, *conn_live;    // [batch]
  float* conn_dist;                    // [batch]
  unsigned long long* best;            // packed (distance bits << 32 | candidate)
  float* box;                          // [22] lo, hi
  float* path;                         // [TGP_MAX_PATH][TGP_STRIDE] device
  float* path_host;                    // pinned
  // shortcut workspace: all pairs of a path of <= 128 waypoints
  float *sc_arm0, *sc_base0, *sc_arm1, *sc_base1;
  uint8_t* sc_valid;
  TgPlanStatus* status;      // host address (mapped)
  TgPlanStatus* status_dev;  // device address of the same memory
  cudaEvent_t ev;
};

__device__ __forceinline__ float tgp_wrap(float a) {  // to (-pi, pi]
  a = fmodf(a + TGP_PI, 2.0f * TGP_PI);
  if (a < 0.f) a += 2.0f * TGP_PI;
  return a - TGP_PI;
}

// squared configuration-space distance with the shortest arc in theta
__device__ __forceinline__ float tgp_dist2(const float* a, const float* b) {
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < TGP_DIM; ++k) {
    float d = b[k] - a[k];
    if (k == 2) d = tgp_wrap(d);
    s = fmaf(d, d, s);
  }
  return s;
}

__device__ __forceinline__ float tgp_halton(unsigned idx, int base) {
  float f = 1.f, r = 0.f;
  const float inv = 1.0f / (float)base;
  while (idx) {
    f *= inv;
    r += f * (float)(idx % base);
    idx /= base;
  }
  return r;
}
__constant__ int TGP_PRIMES[TGP_DIM] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31};

// nearest node of a tree to q (warp-cooperative: lanes stride over the nodes); returns index and squared distance
__device__ __forceinline__ int tgp_nearest(const float* __restrict__ nodes, int n, const float* q, float* d2_out) {
  const unsigned lane = threadIdx.x & 31;
  float best = 3.0e38f;
  int bi = -1;
  for (int i = lane; i < n; i += 32) {
    const float4* row = reinterpret_cast<const float4*>(nodes + (size_t)i * TGP_STRIDE);
    const float4 r0 = __ldg(row), r1 = __ldg(row + 1), r2 = __ldg(row + 2);
    const float v[TGP_STRIDE] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
    const float d2 = tgp_dist2(q, v);
    if (d2 < best) {
      best = d2;
      bi = i;
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob < best || (ob == best && oi >= 0 && (bi < 0 || oi < bi))) {
      best = ob;
      bi = oi;
    }
  }
  *d2_out = best;
  return bi;
}

__device__ __forceinline__ void tgp_store_edge(float* arm0, float* base0, float* arm1, float* base1, int i, const float* a, const float* b) {
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    base0[3 * i + k] = a[k];
    base1[3 * i + k] = b[k];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    arm0[8 * i + k] = a[3 + k];
    arm1[8 * i + k] = b[3 + k];
  }
}

// ------------------------------------------------------------------------------------------------
// RRT-Connect, `batch` extensions per iteration.  One warp per candidate.
// ------------------------------------------------------------------------------------------------
// extend tree `a` towards Halton samples: nearest node, steer at most `range`, emit the edge near -> new
__global__ void __launch_bounds__(128)
k_rrt_extend(TgPlanStatus* st, const float* __restrict__ nodes_a, const int* __restrict__ count, int a, const float* __restrict__ box,
             float range, unsigned halton0, int batch, float* arm0, float* base0, float* arm1, float* base1, int* near_idx) {
  if (st->solved) return;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= batch) return;
  const unsigned lane = threadIdx.x & 31;
  float q[TGP_DIM];
  int free_dim = 0;
#pragma unroll
  for (int k = 0; k < TGP_DIM; ++k) {
    const float lo = box[k], hi = box[TGP_DIM + k];
    q[k] = lo;
    if (hi > lo) {  // fixed dimensions (lo == hi) consume no Halton dimension
      q[k] = lo + (hi - lo) * tgp_halton(halton0 + (unsigned)i + 1u, TGP_PRIMES[free_dim]);
      ++free_dim;
    }
  }
  float d2;
  const int near = tgp_nearest(nodes_a, count[a], q, &d2);
  if (lane == 0) {
    const float* nr = nodes_a + (size_t)near * TGP_STRIDE;
    const float d = sqrtf(d2), s = d > range ? range / d : 1.0f;
    float from[TGP_DIM], to[TGP_DIM];
#pragma unroll
    for (int k = 0; k < TGP_DIM; ++k) {
      from[k] = nr[k];
      float dk = q[k] - nr[k];
      if (k == 2) dk = tgp_wrap(dk);
      to[k] = nr[k] + s * dk;  // theta unwrapped relative to the near node
    }
    tgp_store_edge(arm0, base0, arm1, base1, i, from, to);
    near_idx[i] = near;
  }
}

// append the valid extensions to tree `a`; for each, emit the connection edge new -> nearest node of the other tree
__global__ void __launch_bounds__(128)
k_rrt_commit(TgPlanStatus* st, float* nodes_a, int* parent_a, const float* __restrict__ nodes_b, int* count, int a, int cap, int batch,
             const uint8_t* __restrict__ valid, const int* __restrict__ near_idx, float* arm0, float* base0, float* arm1, float* base1,
             int* conn_a, int* conn_b, int* conn_live, float* conn_dist) {
  if (st->solved) return;
  const int nb = count[1 - a];  // the other tree is not modified in this iteration
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= batch) return;
  const unsigned lane = threadIdx.x & 31;
  float q[TGP_DIM];
#pragma unroll
  for (int k = 0; k < 3; ++k) q[k] = base1[3 * i + k];
#pragma unroll
  for (int k = 0; k < 8; ++k) q[3 + k] = arm1[8 * i + k];
  int slot = -1;
  if (lane == 0 && valid[i]) {
    slot = atomicAdd(&count[a], 1);
    if (slot >= cap) {
      atomicSub(&count[a], 1);
      slot = -1;
    }
  }
  slot = __shfl_sync(0xffffffffu, slot, 0);
  __syncwarp();
  if (slot < 0) {  // nothing added: a degenerate edge (a node of the other tree to itself) keeps the batch dense
    if (lane == 0) {
      tgp_store_edge(arm0, base0, arm1, base1, i, nodes_b, nodes_b);
      conn_live[i] = 0;
    }
    return;
  }
  if (lane < TGP_STRIDE) nodes_a[(size_t)slot * TGP_STRIDE + lane] = lane < TGP_DIM ? q[lane] : 0.f;
  float d2;
  const int nbi = tgp_nearest(nodes_b, nb, q, &d2);
  if (lane == 0) {
    parent_a[slot] = near_idx[i];
    const float* t = nodes_b + (size_t)nbi * TGP_STRIDE;
    float to[TGP_DIM];
#pragma unroll
    for (int k = 0; k < TGP_DIM; ++k) to[k] = k == 2 ? q[2] + tgp_wrap(t[2] - q[2]) : t[k];
    tgp_store_edge(arm0, base0, arm1, base1, i, q, to);
    conn_a[i] = slot;
    conn_b[i] = nbi;
    conn_live[i] = 1;
    conn_dist[i] = sqrtf(d2);
  }
}

// the shortest valid connection (if any) joins the trees
__global__ void __launch_bounds__(256)
k_rrt_finish(TgPlanStatus* st, int a, int batch, const uint8_t* __restrict__ valid, const int* __restrict__ conn_a,
             const int* __restrict__ conn_b, const int* __restrict__ conn_live, const float* __restrict__ conn_dist,
             unsigned long long* best) {
  if (st->solved) return;
  __shared__ unsigned long long s_best;
  if (threadIdx.x == 0) s_best = ~0ull;
  __syncthreads();
  for (int i = threadIdx.x; i < batch; i += blockDim.x)
    if (conn_live[i] && valid[i]) atomicMin(&s_best, ((unsigned long long)__float_as_uint(conn_dist[i]) << 32) | (unsigned)i);
  __syncthreads();
  if (threadIdx.x == 0) {
    st->iterations += 1;
    if (s_best != ~0ull) {
      const int i = (int)(s_best & 0xffffffffu);
      st->tree_of_a = a;
      st->node_a = conn_a[i];
      st->node_b = conn_b[i];
      st->solved = 1;
    }
    *best = s_best;
  }
}

// path start -> goal through the connection, into `path` (theta made continuous along the way)
__global__ void k_rrt_trace(TgPlanStatus* st, const float* nodes0, const int* parent0, const float* nodes1, const int* parent1, float* path) {
  if (!st->solved) {
    st->n_path = 0;
    return;
  }
  const int i0 = st->tree_of_a == 0 ? st->node_a : st->node_b;  // node in the start tree
  const int i1 = st->tree_of_a == 0 ? st->node_b : st->node_a;  // node in the goal tree
  int n0 = 0;
  for (int i = i0; i >= 0; i = parent0[i]) ++n0;
  int n1 = 0;
  for (int i = i1; i >= 0; i = parent1[i]) ++n1;
  if (n0 + n1 > TGP_MAX_PATH) {
    st->n_path = -1;
    return;
  }
  int w = n0 - 1;
  for (int i = i0; i >= 0; i = parent0[i], --w)
    for (int k = 0; k < TGP_STRIDE; ++k) path[(size_t)w * TGP_STRIDE + k] = nodes0[(size_t)i * TGP_STRIDE + k];
  w = n0;
  for (int i = i1; i >= 0; i = parent1[i], ++w)
    for (int k = 0; k < TGP_STRIDE; ++k) path[(size_t)w * TGP_STRIDE + k] = nodes1[(size_t)i * TGP_STRIDE + k];
  for (int j = 1; j < n0 + n1; ++j) {
    float* p = path + (size_t)j * TGP_STRIDE;
    p[2] = p[2 - TGP_STRIDE] + tgp_wrap(p[2] - p[2 - TGP_STRIDE]);
  }
  st->n_path = n0 + n1;
}

// ------------------------------------------------------------------------------------------------
// Shortcutting as ONE batched edge check + a shortest path in the DAG of valid skips: all pairs (i < j) of the
// path's waypoints are validated together; then cost[j] = min over valid (i, j) of cost[i] + |q_j - q_i|.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int tgp_pair_index(int i, int j, int n) { return i * n - (i * (i + 1)) / 2 + (j - i - 1); }

__global__ void __launch_bounds__(128) k_sc_pairs(const float* __restrict__ path, int n, float* arm0, float* base0, float* arm1, float* base1) {
  const int i = blockIdx.x, j = threadIdx.x;
  if (i >= n || j >= n || j <= i) return;
  float a[TGP_DIM], b[TGP_DIM];
#pragma unroll
  for (int k = 0; k < TGP_DIM; ++k) {
    a[k] = path[(size_t)i * TGP_STRIDE + k];
    b[k] = path[(size_t)j * TGP_STRIDE + k];
  }
  b[2] = a[2] + tgp_wrap(b[2] - a[2]);
  tgp_store_edge(arm0, base0, arm1, base1, tgp_pair_index(i, j, n), a, b);
}

// cost of an edge for shortcutting: configuration-space length, times reverse_weight when the base would drive backwards
// (so that simplification does not undo what the base layer's reverse penalty bought)
__device__ __forceinline__ float tgp_reverse_fraction(const float* a, const float* b) {
  // share of the edge (start, middle, end; theta interpolates linearly along it) on which the base moves against its heading
  const float dx = b[0] - a[0], dy = b[1] - a[1];
  if (dx * dx + dy * dy < 1e-10f) return 0.f;
  const float dth = tgp_wrap(b[2] - a[2]);
  int n = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float s, c;
    __sincosf(a[2] + 0.5f * (float)k * dth, &s, &c);
    n += (dx * c + dy * s) < -1e-3f;
  }
  return (float)n * (1.0f / 3.0f);
}
__device__ __forceinline__ float tgp_edge_cost(const float* a, const float* b, float reverse_weight) {
  const float d = sqrtf(tgp_dist2(a, b));
  if (reverse_weight == 1.0f) return d;
  return d * (1.0f + (reverse_weight - 1.0f) * tgp_reverse_fraction(a, b));
}

__global__ void __launch_bounds__(128) k_sc_dp(TgPlanStatus* st, float* path, int n, const uint8_t* __restrict__ valid, float reverse_weight) {
  __shared__ float s_cost[128];
  __shared__ int s_prev[128];
  __shared__ float s_path[128 * TGP_STRIDE];
  __shared__ unsigned long long s_best;
  const int t = threadIdx.x;
  for (int k = t; k < n * TGP_STRIDE; k += blockDim.x) s_path[k] = path[k];
  if (t == 0) {
    s_cost[0] = 0.f;
    s_prev[0] = -1;
  }
  __syncthreads();
  for (int j = 1; j < n; ++j) {
    if (t == 0) s_best = ~0ull;
    __syncthreads();
    if (t < j && (t == j - 1 || valid[tgp_pair_index(t, j, n)])) {  // consecutive waypoints were validated by the planner
      const float c = s_cost[t] + tgp_edge_cost(s_path + t * TGP_STRIDE, s_path + j * TGP_STRIDE, reverse_weight);
      atomicMin(&s_best, ((unsigned long long)__float_as_uint(c) << 32) | (unsigned)t);
    }
    __syncthreads();
    if (t == 0) {
      s_cost[j] = __uint_as_float((unsigned)(s_best >> 32));
      s_prev[j] = (int)(s_best & 0xffffffffu);
    }
    __syncthreads();
  }
  if (t == 0) {
    int m = 0;
    for (int j = n - 1; j >= 0; j = s_prev[j]) ++m;
    int w = m - 1;
    for (int j = n - 1; j >= 0; j = s_prev[j], --w)
      for (int k = 0; k < TGP_STRIDE; ++k) path[(size_t)w * TGP_STRIDE + k] = s_path[j * TGP_STRIDE + k];
    for (int j = 1; j < m; ++j) {
      float* p = path + (size_t)j * TGP_STRIDE;
      p[2] = p[2 - TGP_STRIDE] + tgp_wrap(p[2] - p[2 - TGP_STRIDE]);
    }
    st->n_path = m;
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
#define TGP_SC_MAX 128
#define TGP_SC_PAIRS (TGP_SC_MAX * (TGP_SC_MAX - 1) / 2)

extern "C" int tg_planner_destroy(tg_planner* p) {
  if (!p) return TG_OK;
  cudaSetDevice(p->device);
  for (int t = 0; t < 2; ++t) {
    cudaFree(p->nodes[t]);
    cudaFree(p->parent[t]);
  }
  cudaFree(p->count); cudaFree(p->arm0); cudaFree(p->base0); cudaFree(p->arm1); cudaFree(p->base1); cudaFree(p->valid);
  cudaFree(p->near_idx); cudaFree(p->conn_a); cudaFree(p->conn_b); cudaFree(p->conn_live); cudaFree(p->conn_dist); cudaFree(p->best);
  cudaFree(p->box); cudaFree(p->path); cudaFree(p->sc_arm0); cudaFree(p->sc_base0); cudaFree(p->sc_arm1); cudaFree(p->sc_base1);
  cudaFree(p->sc_valid);
  cudaFreeHost(p->path_host);
  cudaFreeHost(p->status);
  if (p->ev) cudaEventDestroy(p->ev);
  delete p;
  return TG_OK;
}

extern "C" int tg_planner_create(tg_planner** out, tg_env* env, int max_nodes, int batch) {
  TG_CHECK_ARG(out != nullptr && env != nullptr, "tg_planner_create: NULL argument");
  TG_CHECK_ARG(max_nodes >= 64 && max_nodes <= (1 << 22) && batch >= 32 && batch <= (1 << 16), "tg_planner_create: sizes out of range");
  TG_CUDA(cudaSetDevice(env->device));
  tg_planner* p = new tg_planner();
  memset(p, 0, sizeof(*p));
  p->env = env;
  p->device = env->device;
  p->cap = max_nodes;
  p->batch = batch;
#define PL_ALLOC(ptr, bytes)                         \
  do {                                               \
    cudaError_t _e = cudaMalloc(&(ptr), (bytes));    \
    if (_e != cudaSuccess) {                         \
      tg_set_error("tg_planner_create: cudaMalloc failed: %s", cudaGetErrorString(_e)); \
      tg_planner_destroy(p);                         \
      return TG_E_CUDA;                              \
    }                                                \
  } while (0)
  for (int t = 0; t < 2; ++t) {
    PL_ALLOC(p->nodes[t], sizeof(float) * TGP_STRIDE * (size_t)max_nodes);
    PL_ALLOC(p->parent[t], sizeof(int) * (size_t)max_nodes);
  }
  PL_ALLOC(p->count, 2 * sizeof(int));
  PL_ALLOC(p->arm0, sizeof(float) * 8 * (size_t)batch);
  PL_ALLOC(p->arm1, sizeof(float) * 8 * (size_t)batch);
  PL_ALLOC(p->base0, sizeof(float) * 3 * (size_t)batch + 16);
  PL_ALLOC(p->base1, sizeof(float) * 3 * (size_t)batch + 16);
  PL_ALLOC(p->valid, (size_t)batch);
  PL_ALLOC(p->near_idx, sizeof(int) * (size_t)batch);
  PL_ALLOC(p->conn_a, sizeof(int) * (size_t)batch);
  PL_ALLOC(p->conn_b, sizeof(int) * (size_t)batch);
  PL_ALLOC(p->conn_live, sizeof(int) * (size_t)batch);
  PL_ALLOC(p->conn_dist, sizeof(float) * (size_t)batch);
  PL_ALLOC(p->best, sizeof(unsigned long long));
  PL_ALLOC(p->box, sizeof(float) * 2 * TGP_DIM);
  PL_ALLOC(p->path, sizeof(float) * TGP_STRIDE * TGP_MAX_PATH);
  PL_ALLOC(p->sc_arm0, sizeof(float) * 8 * TGP_SC_PAIRS);
  PL_ALLOC(p->sc_arm1, sizeof(float) * 8 * TGP_SC_PAIRS);
  PL_ALLOC(p->sc_base0, sizeof(float) * 3 * TGP_SC_PAIRS + 16);
  PL_ALLOC(p->sc_base1, sizeof(float) * 3 * TGP_SC_PAIRS + 16);
  PL_ALLOC(p->sc_valid, TGP_SC_PAIRS);
#undef PL_ALLOC
  if (cudaMallocHost(&p->path_host, 32768) != cudaSuccess ||  // 256 nodes x 12 floats, or 2048 lattice waypoints x 4 floats
      cudaHostAlloc(&p->status, sizeof(TgPlanStatus), cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer(&p->status_dev, p->status, 0) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev, cudaEventDisableTiming) != cudaSuccess) {
    tg_set_error("tg_planner_create: pinned allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    tg_planner_destroy(p);
    return TG_E_CUDA;
  }
  memset(p->status, 0, sizeof(TgPlanStatus));
  *out = p;
  return TG_OK;
}

static int tgp_fetch_path(tg_planner* p, float* path_out, int max_path, int* n_path, cudaStream_t st) {
  // the ONE device->host copy of a plan: the finished path
  TG_CUDA(cudaStreamSynchronize(st));
  const int n = p->status->n_path;
  *n_path = n;
  if (n <= 0) return TG_OK;
  if (n > max_path) {
    tg_set_error("planner: path of %d waypoints exceeds the caller's buffer (%d)", n, max_path);
    return TG_E_CAPACITY;
  }
  TG_CUDA(cudaMemcpyAsync(p->path_host, p->path, sizeof(float) * TGP_STRIDE * n, cudaMemcpyDeviceToHost, st));
  TG_CUDA(cudaStreamSynchronize(st));
  for (int i = 0; i < n; ++i)
    for (int k = 0; k < TGP_DIM; ++k) path_out[i * TGP_DIM + k] = p->path_host[i * TGP_STRIDE + k];
  return TG_OK;
}

// Batched RRT-Connect between two configurations.  start / goal / lo / hi: HOST, 11 floats [x, y, theta, q0..q7];
// a dimension with lo == hi is not sampled (arm-only planning fixes the base).  Every iteration extends one tree by
// up to `batch` validated edges and tries to connect every new node to the other tree.  The host enqueues `poll`
// iterations, then waits on an event and reads the status word (mapped memory): no copies inside the loop.
// Returns TG_OK with *n_path = 0 when no path was found within max_iters / timeout_s.
extern "C" int tg_planner_rrt_connect(tg_planner* p, const float* start, const float* goal, const float* lo, const float* hi, float range,
                                      int max_iters, uint32_t halton_skip, int rake, float resolution, double timeout_s, float* path_out,
                                      int max_path, int* n_path, int* iterations, void* stream_) {
  TG_CHECK_ARG(p && start && goal && lo && hi && path_out && n_path && iterations, "tg_planner_rrt_connect: NULL argument");
  TG_CHECK_ARG(range > 0.f && max_iters >= 1 && max_path >= 2, "tg_planner_rrt_connect: bad parameter");
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(p->device));
  *n_path = 0;
  *iterations = 0;
  float row[2][TGP_STRIDE] = {{0}}, box[2 * TGP_DIM];
  for (int k = 0; k < TGP_DIM; ++k) {
    row[0][k] = start[k];
    row[1][k] = goal[k];
    box[k] = lo[k];
    box[TGP_DIM + k] = hi[k];
  }
  const int one[2] = {1, 1}, minus1 = -1;
  memset(p->status, 0, sizeof(TgPlanStatus));
  // setup (before the loop): roots, sampling box
  for (int t = 0; t < 2; ++t) {
    TG_CUDA(cudaMemcpyAsync(p->nodes[t], row[t], sizeof(float) * TGP_STRIDE, cudaMemcpyHostToDevice, st));
    TG_CUDA(cudaMemcpyAsync(p->parent[t], &minus1, sizeof(int), cudaMemcpyHostToDevice, st));
  }
  TG_CUDA(cudaMemcpyAsync(p->count, one, sizeof(one), cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaMemcpyAsync(p->box, box, sizeof(box), cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaStreamSynchronize(st));  // the staging arrays above live on this stack frame
  const int warps_per_block = 4, blocks = tg_div_up(p->batch, warps_per_block);
  const auto t0 = std::chrono::steady_clock::now();
  const int poll = 4;
  int it = 0;
  while (it < max_iters && !p->status->solved) {
    for (int k = 0; k < poll && it < max_iters; ++k, ++it) {
      const int a = it & 1;
      k_rrt_extend<<<blocks, 32 * warps_per_block, 0, st>>>(p->status_dev, p->nodes[a], p->count, a, p->box, range,
                                                            halton_skip + (unsigned)it * (unsigned)p->batch, p->batch, p->arm0, p->base0,
                                                            p->arm1, p->base1, p->near_idx);
      int rc = tg_check_edges(p->env, p->arm0, p->base0, p->arm1, p->base1, p->batch, rake, resolution, p->valid, nullptr, st);
      if (rc) return rc;
      k_rrt_commit<<<blocks, 32 * warps_per_block, 0, st>>>(p->status_dev, p->nodes[a], p->parent[a], p->nodes[1 - a], p->count, a, p->cap,
                                                            p->batch, p->valid, p->near_idx, p->arm0, p->base0, p->arm1, p->base1, p->conn_a,
                                                            p->conn_b, p->conn_live, p->conn_dist);
      rc = tg_check_edges(p->env, p->arm0, p->base0, p->arm1, p->base1, p->batch, rake, resolution, p->valid, nullptr, st);
      if (rc) return rc;
      k_rrt_finish<<<1, 256, 0, st>>>(p->status_dev, a, p->batch, p->valid, p->conn_a, p->conn_b, p->conn_live, p->conn_dist, p->best);
    }
    TG_CUDA(cudaEventRecord(p->ev, st));
    TG_CUDA(cudaEventSynchronize(p->ev));
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) break;
  }
  TG_CUDA(cudaGetLastError());
  *iterations = p->status->iterations;
  if (!p->status->solved) return TG_OK;
  k_rrt_trace<<<1, 1, 0, st>>>(p->status_dev, p->nodes[0], p->parent[0], p->nodes[1], p->parent[1], p->path);
  return tgp_fetch_path(p, path_out, max_path, n_path, st);
}

// Shortcut a path (HOST in/out, n x 11, n <= 128) by validating ALL waypoint pairs in one tg_check_edges call and
// taking the shortest path through the valid skips.  Consecutive waypoints are taken as valid (the planner's edges).
extern "C" int tg_planner_shortcut(tg_planner* p, float* path, int* n_io, int rake, float resolution, float reverse_weight, void* stream_) {
  TG_CHECK_ARG(p && path && n_io, "tg_planner_shortcut: NULL argument");
  const int n = *n_io;
  if (n <= 2) return TG_OK;
  TG_CHECK_ARG(n <= TGP_SC_MAX, "tg_planner_shortcut: at most 128 waypoints");
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(p->device));
  for (int i = 0; i < n; ++i) {
    for (int k = 0; k < TGP_DIM; ++k) p->path_host[i * TGP_STRIDE + k] = path[i * TGP_DIM + k];
    p->path_host[i * TGP_STRIDE + TGP_DIM] = 0.f;
  }
  TG_CUDA(cudaMemcpyAsync(p->path, p->path_host, sizeof(float) * TGP_STRIDE * n, cudaMemcpyHostToDevice, st));
  k_sc_pairs<<<n, 128, 0, st>>>(p->path, n, p->sc_arm0, p->sc_base0, p->sc_arm1, p->sc_base1);
  const int pairs = n * (n - 1) / 2;
  const int rc = tg_check_edges(p->env, p->sc_arm0, p->sc_base0, p->sc_arm1, p->sc_base1, pairs, rake, resolution, p->sc_valid, nullptr, st);
  if (rc) return rc;
  k_sc_dp<<<1, 128, 0, st>>>(p->status_dev, p->path, n, p->sc_valid, reverse_weight >= 1.0f ? reverse_weight : 1.0f);
  TG_CUDA(cudaGetLastError());
  return tgp_fetch_path(p, path, n, n_io, st);
}

// ------------------------------------------------------------------------------------------------
// Base layer of the multilayer planner: SE(2) lattice search (the reference configures VAMP's hybrid A* with
// reverse_penalty = 50, planning/fetch/algorithms.py:93-96).  States (ix, iy, k): a square grid of `res` metres and 16
// headings whose direction vectors are lattice vectors, so every move lands exactly on a node.  Moves: forward along
// the heading, backward (cost x reverse_penalty), turn in place to a neighbouring heading.  Cost-to-go from the goal
// by parallel relaxation (Bellman-Ford sweeps over all nodes, one thread per node, until a sweep changes nothing) -
// data-parallel instead of a priority queue.  Node validity comes from ONE tg_check_configs call over the lattice.
// ------------------------------------------------------------------------------------------------
__constant__ int TGL_DX[16] = {1, 2, 1, 1, 0, -1, -1, -2, -1, -2, -1, -1, 0, 1, 1, 2};
__constant__ int TGL_DY[16] = {0, 1, 1, 2, 1, 2, 1, 1, 0, -1, -1, -2, -1, -2, -1, -1};

struct TgLattice {
  int nx, ny;
  float x0, y0, res;
  float reverse_penalty, turn_cost;  // turn_cost: metres per radian of turning in place
};

__device__ __forceinline__ float tgl_heading(int k) { return atan2f((float)TGL_DY[k], (float)TGL_DX[k]); }

// lattice node -> whole-body configuration with the given arm (for the validity check)
__global__ void __launch_bounds__(256) k_lat_configs(TgLattice L, const float* __restrict__ arm8, float* arm, float* base) {
  const int n = L.nx * L.ny * 16, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = i & 15, c = i >> 4, ix = c % L.nx, iy = c / L.nx;
  base[3 * (size_t)i] = L.x0 + L.res * (float)ix;
  base[3 * (size_t)i + 1] = L.y0 + L.res * (float)iy;
  base[3 * (size_t)i + 2] = tgl_heading(k);
#pragma unroll
  for (int j = 0; j < 8; ++j) arm[8 * (size_t)i + j] = arm8[j];
}

// successor of node (ix, iy, k) under move m: 0 forward, 1 backward, 2 turn left, 3 turn right; -1 if outside
__device__ __forceinline__ int tgl_succ(const TgLattice& L, int ix, int iy, int k, int m, float* cost) {
  int jx = ix, jy = iy, jk = k;
  const float step = L.res * sqrtf((float)(TGL_DX[k] * TGL_DX[k] + TGL_DY[k] * TGL_DY[k]));
  if (m == 0) { jx += TGL_DX[k]; jy += TGL_DY[k]; *cost = step; }
  else if (m == 1) { jx -= TGL_DX[k]; jy -= TGL_DY[k]; *cost = step * L.reverse_penalty; }
  else {
    jk = (k + (m == 2 ? 1 : 15)) & 15;
    float da = fabsf(tgl_heading(jk) - tgl_heading(k));
    if (da > TGP_PI) da = 2.0f * TGP_PI - da;
    *cost = L.turn_cost * da;
  }
  if (jx < 0 || jx >= L.nx || jy < 0 || jy >= L.ny) return -1;
  return ((jy * L.nx + jx) << 4) | jk;
}

// one relaxation sweep: dist[v] = min over moves of cost + dist[succ] (valid nodes, unblocked moves only)
__global__ void __launch_bounds__(256)
k_lat_relax(TgPlanStatus* st, TgLattice L, const uint8_t* __restrict__ node_valid, const uint8_t* __restrict__ blocked, float* dist) {
  const int n = L.nx * L.ny * 16, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !node_valid[i]) return;
  const int k = i & 15, c = i >> 4, ix = c % L.nx, iy = c / L.nx;
  float best = dist[i];
  bool better = false;
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    float cost;
    const int j = tgl_succ(L, ix, iy, k, m, &cost);
    if (j < 0 || !node_valid[j] || (blocked[i] >> m) & 1) continue;
    const float d = cost + dist[j];  // racing with the successor's own update is harmless: values only decrease
    if (d < best) {
      best = d;
      better = true;
    }
  }
  if (better) {
    dist[i] = best;
    st->changed = 1;
  }
}

// waypoints: the continuous start pose, the cheapest lattice moves from the snapped start node to the snapped goal node,
// the continuous goal pose; theta continuous along the list.  wp rows: x, y, theta, move taken FROM this waypoint.
__global__ void k_lat_walk(TgPlanStatus* st, TgLattice L, const uint8_t* __restrict__ node_valid, const uint8_t* __restrict__ blocked,
                           const float* __restrict__ dist, const int* __restrict__ nodes2 /* start, goal */, float sx, float sy, float sth,
                           float gx, float gy, float gth, float* wp, int* wp_node, int max_wp) {
  st->n_path = 0;
  const int start_node = nodes2[0], goal_node = nodes2[1];
  if (start_node < 0 || goal_node < 0 || !(dist[start_node] < 1.0e38f)) return;
  wp[0] = sx; wp[1] = sy; wp[2] = sth; wp[3] = -1.f;
  wp_node[0] = -1;
  int v = start_node, n = 1;
  float th_prev = sth;
  bool reached = false;
  while (n < max_wp - 1) {
    const int k = v & 15, c = v >> 4, ix = c % L.nx, iy = c / L.nx;
    const float th = th_prev + tgp_wrap(tgl_heading(k) - th_prev);
    th_prev = th;
    wp[4 * n] = L.x0 + L.res * (float)ix;
    wp[4 * n + 1] = L.y0 + L.res * (float)iy;
    wp[4 * n + 2] = th;
    wp[4 * n + 3] = -1.f;
    wp_node[n] = v;
    ++n;
    if (v == goal_node) {
      reached = true;
      break;
    }
    float best = 3.0e38f;
    int bj = -1, bm = 0;
    for (int m = 0; m < 4; ++m) {
      float cost;
      const int j = tgl_succ(L, ix, iy, k, m, &cost);
      if (j < 0 || !node_valid[j] || (blocked[v] >> m) & 1) continue;
      if (cost + dist[j] < best) {
        best = cost + dist[j];
        bj = j;
        bm = m;
      }
    }
    if (bj < 0) return;
    wp[4 * (n - 1) + 3] = (float)bm;
    v = bj;
  }
  if (!reached) return;
  wp[4 * n] = gx; wp[4 * n + 1] = gy; wp[4 * n + 2] = th_prev + tgp_wrap(gth - th_prev); wp[4 * n + 3] = -1.f;
  wp_node[n] = -1;
  st->n_path = n + 1;
}

// whole-body edges along the base path (arm fixed) for validation
__global__ void __launch_bounds__(128)
k_lat_edges(const float* __restrict__ wp, int n, const float* __restrict__ arm8, float* arm0, float* base0, float* arm1, float* base1) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    base0[3 * i + k] = wp[4 * i + k];
    base1[3 * i + k] = wp[4 * (i + 1) + k];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) arm0[8 * i + k] = arm1[8 * i + k] = arm8[k];
}

// block the moves whose swept edge collided (the next relaxation routes around them)
__global__ void __launch_bounds__(128)
k_lat_block(TgPlanStatus* st, const float* __restrict__ wp, const int* __restrict__ wp_node, int n, const uint8_t* __restrict__ valid,
            uint8_t* blocked, uint8_t* node_valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1 || valid[i]) return;
  st->aux = 1;  // at least one edge of the path was refuted
  if (wp_node[i] < 0) {  // start pose -> snapped start node collided: snap elsewhere next round
    node_valid[wp_node[i + 1]] = 0;
    return;
  }
  if (wp_node[i + 1] < 0) {  // snapped goal node -> goal pose
    node_valid[wp_node[i]] = 0;
    return;
  }
  const int m = (int)wp[4 * i + 3];
  atomicOr(reinterpret_cast<unsigned*>(blocked + (wp_node[i] & ~3)), (1u << m) << (8 * (wp_node[i] & 3)));
}

// nearest VALID lattice node to a continuous pose (within +-3 cells, any heading): out[0] = node or -1
__global__ void __launch_bounds__(256) k_lat_snap(TgLattice L, const uint8_t* __restrict__ node_valid, float x, float y, float th, int* out) {
  __shared__ unsigned long long s_best;
  if (threadIdx.x == 0) s_best = ~0ull;
  __syncthreads();
  const int cx = (int)rintf((x - L.x0) / L.res), cy = (int)rintf((y - L.y0) / L.res);
  for (int t = threadIdx.x; t < 7 * 7 * 16; t += blockDim.x) {
    const int k = t & 15, c = t >> 4, ix = cx + c % 7 - 3, iy = cy + c / 7 - 3;
    if (ix < 0 || ix >= L.nx || iy < 0 || iy >= L.ny) continue;
    const int v = ((iy * L.nx + ix) << 4) | k;
    if (!node_valid[v]) continue;
    const float dx = L.x0 + L.res * ix - x, dy = L.y0 + L.res * iy - y, da = tgp_wrap(tgl_heading(k) - th);
    const float d = sqrtf(dx * dx + dy * dy) + L.turn_cost * fabsf(da);
    atomicMin(&s_best, ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)v);
  }
  __syncthreads();
  if (threadIdx.x == 0) out[0] = s_best == ~0ull ? -1 : (int)(s_best & 0xffffffffu);
}
__global__ void k_lat_init_dev(TgLattice L, float* dist, const int* goal_node) {
  const int n = L.nx * L.ny * 16, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dist[i] = i == goal_node[0] ? 0.f : 3.0e38f;
}
__global__ void k_plan_clear(TgPlanStatus* st, int what) {
  if (what & 1) st->changed = 0;
  if (what & 2) st->aux = 0;
}

struct TgLatticeBuf {  // grown on demand, owned by the planner handle (file-static map keeps tg_planner POD-simple)
  size_t nodes;
  float *dist, *arm, *base, *wp, *arm8;
  uint8_t *node_valid, *blocked, *evalid;
  int *wp_node, *snap;
  float *e_arm0, *e_base0, *e_arm1, *e_base1;
};
#define TGL_MAX_WP 2048

static void tgl_free(TgLatticeBuf* b) {
  cudaFree(b->dist); cudaFree(b->arm); cudaFree(b->base); cudaFree(b->wp); cudaFree(b->arm8); cudaFree(b->node_valid);
  cudaFree(b->blocked); cudaFree(b->evalid); cudaFree(b->wp_node); cudaFree(b->snap); cudaFree(b->e_arm0); cudaFree(b->e_base0);
  cudaFree(b->e_arm1); cudaFree(b->e_base1);
  memset(b, 0, sizeof(*b));
}

// Base path on the SE(2) lattice with the arm held at `arm8`.  start / goal: HOST [x, y, theta]; bounds: HOST
// {x0, x1, y0, y1}.  Lazy edge validation: the cheapest lattice path is validated as whole-body edges in one
// tg_check_edges call; refuted moves are blocked and the search repeats (at most max_rounds times).
// path_out: HOST max_path x 3 (start pose, lattice waypoints, goal pose; theta continuous).  stats3 (HOST): relaxation
// sweeps, validation rounds, lattice nodes.  *n_path = 0 when the base layer finds no path.
extern "C" int tg_planner_base_path(tg_planner* p, const float* start, const float* goal, const float* arm8, const float* bounds4, float res,
                                    float reverse_penalty, float turn_cost, int rake, float resolution, int max_rounds, float* path_out,
                                    int max_path, int* n_path, double* stats3, void* stream_) {
  TG_CHECK_ARG(p && start && goal && arm8 && bounds4 && path_out && n_path, "tg_planner_base_path: NULL argument");
  TG_CHECK_ARG(res > 0.f && bounds4[1] > bounds4[0] && bounds4[3] > bounds4[2] && max_rounds >= 1, "tg_planner_base_path: bad parameter");
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(p->device));
  *n_path = 0;
  TgLattice L;
  L.x0 = bounds4[0]; L.y0 = bounds4[2]; L.res = res;
  L.nx = (int)floorf((bounds4[1] - bounds4[0]) / res) + 1;
  L.ny = (int)floorf((bounds4[3] - bounds4[2]) / res) + 1;
  L.reverse_penalty = reverse_penalty;
  L.turn_cost = turn_cost;
  const size_t n = (size_t)L.nx * L.ny * 16;
  TG_CHECK_ARG(n <= ((size_t)1 << 24), "tg_planner_base_path: lattice too large (coarsen the resolution)");
  static TgLatticeBuf buf;  // one planner at a time per process (the reference plans single-threaded)
  if (buf.nodes < n) {
    tgl_free(&buf);
    const size_t cap = n + n / 4 + 1024;
    TG_CUDA(cudaMalloc(&buf.dist, 4 * cap));
    TG_CUDA(cudaMalloc(&buf.arm, 32 * cap));
    TG_CUDA(cudaMalloc(&buf.base, 12 * cap + 16));
    TG_CUDA(cudaMalloc(&buf.node_valid, cap));
    TG_CUDA(cudaMalloc(&buf.blocked, cap + 4));
    TG_CUDA(cudaMalloc(&buf.wp, 16 * TGL_MAX_WP));
    TG_CUDA(cudaMalloc(&buf.wp_node, 4 * TGL_MAX_WP));
    TG_CUDA(cudaMalloc(&buf.snap, 8));
    TG_CUDA(cudaMalloc(&buf.arm8, 32));
    TG_CUDA(cudaMalloc(&buf.evalid, TGL_MAX_WP));
    TG_CUDA(cudaMalloc(&buf.e_arm0, 32 * TGL_MAX_WP));
    TG_CUDA(cudaMalloc(&buf.e_arm1, 32 * TGL_MAX_WP));
    TG_CUDA(cudaMalloc(&buf.e_base0, 12 * TGL_MAX_WP + 16));
    TG_CUDA(cudaMalloc(&buf.e_base1, 12 * TGL_MAX_WP + 16));
    buf.nodes = cap;
  }
  memset(p->status, 0, sizeof(TgPlanStatus));
  TG_CUDA(cudaMemcpyAsync(buf.arm8, arm8, 32, cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaStreamSynchronize(st));
  const int nb = tg_div_up((int64_t)n, 256);
  k_lat_configs<<<nb, 256, 0, st>>>(L, buf.arm8, buf.arm, buf.base);
  int rc = tg_check_configs(p->env, buf.arm, buf.base, (int64_t)n, buf.node_valid, nullptr, st);
  if (rc) return rc;
  TG_CUDA(cudaMemsetAsync(buf.blocked, 0, n + 4, st));
  int sweeps = 0, rounds = 0;
  for (; rounds < max_rounds; ++rounds) {
    k_lat_snap<<<1, 256, 0, st>>>(L, buf.node_valid, start[0], start[1], start[2], buf.snap);
    k_lat_snap<<<1, 256, 0, st>>>(L, buf.node_valid, goal[0], goal[1], goal[2], buf.snap + 1);
    k_lat_init_dev<<<nb, 256, 0, st>>>(L, buf.dist, buf.snap + 1);
    for (;;) {  // relaxation to the fixed point, polled every 16 sweeps
      k_plan_clear<<<1, 1, 0, st>>>(p->status_dev, 1);
      for (int k = 0; k < 16; ++k) k_lat_relax<<<nb, 256, 0, st>>>(p->status_dev, L, buf.node_valid, buf.blocked, buf.dist);
      sweeps += 16;
      TG_CUDA(cudaEventRecord(p->ev, st));
      TG_CUDA(cudaEventSynchronize(p->ev));
      if (!p->status->changed || sweeps > 16384) break;
    }
    k_plan_clear<<<1, 1, 0, st>>>(p->status_dev, 2);
    k_lat_walk<<<1, 1, 0, st>>>(p->status_dev, L, buf.node_valid, buf.blocked, buf.dist, buf.snap, start[0], start[1], start[2], goal[0],
                                goal[1], goal[2], buf.wp, buf.wp_node, TGL_MAX_WP);
    TG_CUDA(cudaEventRecord(p->ev, st));
    TG_CUDA(cudaEventSynchronize(p->ev));
    const int nw = p->status->n_path;
    if (nw < 2) break;  // unreachable
    k_lat_edges<<<tg_div_up(nw - 1, 128), 128, 0, st>>>(buf.wp, nw, buf.arm8, buf.e_arm0, buf.e_base0, buf.e_arm1, buf.e_base1);
    rc = tg_check_edges(p->env, buf.e_arm0, buf.e_base0, buf.e_arm1, buf.e_base1, nw - 1, rake, resolution, buf.evalid, nullptr, st);
    if (rc) return rc;
    k_lat_block<<<tg_div_up(nw - 1, 128), 128, 0, st>>>(p->status_dev, buf.wp, buf.wp_node, nw, buf.evalid, buf.blocked, buf.node_valid);
    TG_CUDA(cudaEventRecord(p->ev, st));
    TG_CUDA(cudaEventSynchronize(p->ev));
    if (!p->status->aux) {  // every edge of the path is collision-free: the one device->host copy, the waypoints
      if (nw > max_path) {
        tg_set_error("tg_planner_base_path: %d waypoints exceed the caller's buffer (%d)", nw, max_path);
        return TG_E_CAPACITY;
      }
      TG_CUDA(cudaMemcpyAsync(p->path_host, buf.wp, 16 * (size_t)nw, cudaMemcpyDeviceToHost, st));
      TG_CUDA(cudaStreamSynchronize(st));
      for (int i = 0; i < nw; ++i)
        for (int k = 0; k < 3; ++k) path_out[3 * i + k] = p->path_host[4 * i + k];
      *n_path = nw;
      ++rounds;
      break;
    }
  }
  if (stats3) {
    stats3[0] = sweeps;
    stats3[1] = rounds;
    stats3[2] = (double)n;
  }
  TG_CUDA(cudaGetLastError());
  return TG_OK;
}

// ------------------------------------------------------------------------------------------------
// fcit_wb: a FULLY CONNECTED, fully evaluated roadmap (the premise of FCIT*, planning/fetch/algorithms.py:155-232 with
// batch_size 4096 / max_samples 8192 / reverse_weight 10: edge checks are cheap enough to connect everything): Halton
// samples inside the cloud's XY bounds are validated in one tg_check_configs call; EVERY pair of valid samples closer
// than max_edge becomes an edge and all of them (10^5 - 10^6) are validated in ONE tg_check_edges call - a few ms on this
// machine, where a lazy search needed hundreds of serial rounds in clutter; the cost-to-go over the valid graph is relaxed
// in parallel (one thread per node, all other nodes streamed through shared memory - a min-plus matrix-vector product
// per sweep).  A further batch of samples is added (only the new pairs are validated) while the start is unreachable.
// ------------------------------------------------------------------------------------------------
// Three of four samples keep the arm on the start configuration, on the goal configuration or on the segment between
// them (the base is always sampled): a whole-body query mostly moves the BASE, and a roadmap of uniformly random arm
// postures would make every edge a 5 rad arm swing through the clutter.
__global__ void __launch_bounds__(256)
k_fcit_samples(const float* __restrict__ box, const float* __restrict__ ends /* nodes 0 (start) and 1 (goal) */, unsigned halton0, int n,
               float* arm, float* base) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float q[TGP_DIM];
  int free_dim = 0;
#pragma unroll
  for (int k = 0; k < TGP_DIM; ++k) {
    const float lo = box[k], hi = box[TGP_DIM + k];
    q[k] = lo;
    if (hi > lo) {
      q[k] = lo + (hi - lo) * tgp_halton(halton0 + (unsigned)i + 1u, TGP_PRIMES[free_dim]);
      ++free_dim;
    }
  }
  const unsigned mode = (halton0 + (unsigned)i) & 3u;
  if (mode) {
    const float t = mode == 1 ? 0.f : (mode == 2 ? 1.f : tgp_halton(halton0 + (unsigned)i + 1u, 37));
#pragma unroll
    for (int k = 3; k < TGP_DIM; ++k) q[k] = ends[k] + t * (ends[TGP_STRIDE + k] - ends[k]);
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) base[3 * (size_t)i + k] = q[k];
#pragma unroll
  for (int k = 0; k < 8; ++k) arm[8 * (size_t)i + k] = q[3 + k];
}

__global__ void __launch_bounds__(256)
k_fcit_append(const uint8_t* __restrict__ valid, const float* __restrict__ arm, const float* __restrict__ base, int n, float* nodes, int* count,
              int cap) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !valid[i]) return;
  const int slot = atomicAdd(count, 1);
  if (slot >= cap) {
    atomicSub(count, 1);
    return;
  }
  float* r = nodes + (size_t)slot * TGP_STRIDE;
#pragma unroll
  for (int k = 0; k < 3; ++k) r[k] = base[3 * (size_t)i + k];
#pragma unroll
  for (int k = 0; k < 8; ++k) r[3 + k] = arm[8 * (size_t)i + k];
  r[11] = 0.f;
}

__global__ void __launch_bounds__(256) k_fcit_init(float* dist, int* next, int cap) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cap) {
    dist[i] = i == 1 ? 0.f : 3.0e38f;  // node 1 is the goal
    next[i] = -1;
  }
}

// cost of driving v -> u: configuration-space length, times reverse_weight when the base would move backwards
__device__ __forceinline__ bool tgf_cost(const float* v, const float* u, float reverse_weight, float max_edge, float* w) {
  const float d = sqrtf(tgp_dist2(v, u));
  if (d > max_edge) return false;  // the graph is complete up to this edge length
  *w = reverse_weight == 1.0f ? d : d * (1.0f + (reverse_weight - 1.0f) * tgp_reverse_fraction(v, u));
  return true;
}

__global__ void __launch_bounds__(256)
k_fcit_relax(TgPlanStatus* st, const float* __restrict__ nodes, const int* __restrict__ count, const unsigned* __restrict__ ok, int words,
             float reverse_weight, float max_edge, float* dist, int* next) {
  __shared__ float s_nodes[256 * TGP_STRIDE];
  __shared__ float s_dist[256];
  const int n = *count, v = blockIdx.x * blockDim.x + threadIdx.x;
  if (blockIdx.x * blockDim.x >= n) return;
  float me[TGP_STRIDE];
  if (v < n) {
#pragma unroll
    for (int k = 0; k < TGP_STRIDE; ++k) me[k] = nodes[(size_t)v * TGP_STRIDE + k];
  }
  float best = v < n ? dist[v] : 0.f;
  int bn = -1;
  for (int u0 = 0; u0 < n; u0 += 256) {
    __syncthreads();
    for (int k = threadIdx.x; k < 256 * TGP_STRIDE; k += 256) {
      const int u = u0 + k / TGP_STRIDE;
      s_nodes[k] = u < n ? nodes[(size_t)u0 * TGP_STRIDE + k] : 0.f;
    }
    s_dist[threadIdx.x] = u0 + (int)threadIdx.x < n ? dist[u0 + threadIdx.x] : 3.0e38f;
    __syncthreads();
    if (v >= n || v == 1) continue;
    const int m = min(256, n - u0);
    for (int j = 0; j < m; ++j) {
      const float du = s_dist[j];
      if (!(du < best)) continue;  // cannot improve (also skips unreached nodes)
      const int u = u0 + j;
      if (u == v || !((ok[(size_t)v * words + (u >> 5)] >> (u & 31)) & 1u)) continue;  // only validated edges
      float w;
      if (!tgf_cost(me, s_nodes + j * TGP_STRIDE, reverse_weight, max_edge, &w)) continue;
      if (w + du < best) {
        best = w + du;
        bn = u;
      }
    }
  }
  if (v < n && bn >= 0) {
    dist[v] = best;
    next[v] = bn;
    st->changed = 1;
  }
}

// all pairs (i < j, j >= j0) of roadmap nodes closer than max_edge -> edge list (theta of the target unwrapped)
__global__ void __launch_bounds__(256)
k_fcit_pairs(const float* __restrict__ nodes, const int* __restrict__ count, int j0, float max_edge, float* arm0, float* base0, float* arm1,
             float* base1, int2* pair, int* n_edges, int cap) {
  const int n = *count, j = j0 + blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= n || i >= j) return;
  float a[TGP_STRIDE], b[TGP_STRIDE];
#pragma unroll
  for (int k = 0; k < TGP_STRIDE; ++k) {
    a[k] = nodes[(size_t)i * TGP_STRIDE + k];
    b[k] = nodes[(size_t)j * TGP_STRIDE + k];
  }
  if (tgp_dist2(a, b) > max_edge * max_edge) return;
  const int e = atomicAdd(n_edges, 1);
  if (e >= cap) return;  // counted, not stored: the host sees the overflow
  b[2] = a[2] + tgp_wrap(b[2] - a[2]);
  tgp_store_edge(arm0, base0, arm1, base1, e, a, b);
  pair[e] = make_int2(i, j);
}

__global__ void __launch_bounds__(256)
k_fcit_mark(const uint8_t* __restrict__ valid, const int2* __restrict__ pair, int E, unsigned* ok, int words) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E || !valid[e]) return;
  const int2 p = pair[e];
  atomicOr(&ok[(size_t)p.x * words + (p.y >> 5)], 1u << (p.y & 31));
  atomicOr(&ok[(size_t)p.y * words + (p.x >> 5)], 1u << (p.x & 31));
}

// path node 0 (start) -> node 1 (goal) along next[] (every edge of it was validated)
__global__ void k_fcit_path(TgPlanStatus* st, const float* __restrict__ nodes, const float* __restrict__ dist, const int* __restrict__ next,
                            float* path, int max_n) {
  st->n_path = 0;
  if (!(dist[0] < 1.0e38f)) return;
  int v = 0, n = 0;
  while (n < max_n) {
    for (int k = 0; k < TGP_STRIDE; ++k) path[(size_t)n * TGP_STRIDE + k] = nodes[(size_t)v * TGP_STRIDE + k];
    if (n > 0) {
      float* p = path + (size_t)n * TGP_STRIDE;
      p[2] = p[2 - TGP_STRIDE] + tgp_wrap(p[2] - p[2 - TGP_STRIDE]);
    }
    ++n;
    if (v == 1) {
      st->n_path = n;
      return;
    }
    v = next[v];
    if (v < 0) return;
  }
}

__global__ void k_plan_publish(TgPlanStatus* st, const int* value) { st->aux = *value; }

struct TgFcitBuf {
  int cap, batch, ecap;
  float *nodes, *dist, *arm, *base;
  int *next, *count, *n_edges;
  unsigned* ok;
  uint8_t *valid, *evalid;
  float *e_arm0, *e_base0, *e_arm1, *e_base1;
  int2* pair;
};
#define TGF_MAX_EDGES (1 << 21)

// start / goal: HOST 11-vectors; lo / hi: HOST sampling box (XY from the cloud bounds, theta, joint limits).
// stats4 (HOST): samples drawn, roadmap nodes, validated edges, relaxation sweeps.  *n_path = 0: no path within the budgets.
extern "C" int tg_planner_fcit(tg_planner* p, const float* start, const float* goal, const float* lo, const float* hi, int batch_size,
                               int max_samples, int max_rounds, float reverse_weight, float max_edge, uint32_t halton_skip, int rake,
                               float resolution, double timeout_s, float* path_out, int max_path, int* n_path, double* stats4, void* stream_) {
  TG_CHECK_ARG(p && start && goal && lo && hi && path_out && n_path, "tg_planner_fcit: NULL argument");
  TG_CHECK_ARG(batch_size >= 32 && max_samples >= batch_size && max_samples <= (1 << 16) && max_rounds >= 1 && max_edge > 0.f,
               "tg_planner_fcit: bad parameter");
  cudaStream_t st = (cudaStream_t)stream_;
  TG_CUDA(cudaSetDevice(p->device));
  *n_path = 0;
  static TgFcitBuf buf;
  const int cap = max_samples + 2, words = (cap + 31) / 32;
  if (buf.cap < cap || buf.batch < batch_size) {
    cudaFree(buf.nodes); cudaFree(buf.dist); cudaFree(buf.arm); cudaFree(buf.base); cudaFree(buf.next); cudaFree(buf.count);
    cudaFree(buf.n_edges); cudaFree(buf.ok); cudaFree(buf.valid); cudaFree(buf.evalid); cudaFree(buf.e_arm0); cudaFree(buf.e_base0);
    cudaFree(buf.e_arm1); cudaFree(buf.e_base1); cudaFree(buf.pair);
    memset(&buf, 0, sizeof(buf));
    TG_CUDA(cudaMalloc(&buf.nodes, sizeof(float) * TGP_STRIDE * (size_t)cap));
    TG_CUDA(cudaMalloc(&buf.dist, 4 * (size_t)cap));
    TG_CUDA(cudaMalloc(&buf.next, 4 * (size_t)cap));
    TG_CUDA(cudaMalloc(&buf.count, 4));
    TG_CUDA(cudaMalloc(&buf.n_edges, 4));
    TG_CUDA(cudaMalloc(&buf.ok, 4 * (size_t)cap * words));
    TG_CUDA(cudaMalloc(&buf.arm, 32 * (size_t)batch_size));
    TG_CUDA(cudaMalloc(&buf.base, 12 * (size_t)batch_size + 16));
    TG_CUDA(cudaMalloc(&buf.valid, (size_t)batch_size));
    TG_CUDA(cudaMalloc(&buf.evalid, TGF_MAX_EDGES));
    TG_CUDA(cudaMalloc(&buf.e_arm0, 32 * (size_t)TGF_MAX_EDGES));
    TG_CUDA(cudaMalloc(&buf.e_arm1, 32 * (size_t)TGF_MAX_EDGES));
    TG_CUDA(cudaMalloc(&buf.e_base0, 12 * (size_t)TGF_MAX_EDGES + 16));
    TG_CUDA(cudaMalloc(&buf.e_base1, 12 * (size_t)TGF_MAX_EDGES + 16));
    TG_CUDA(cudaMalloc(&buf.pair, 8 * (size_t)TGF_MAX_EDGES));
    buf.cap = cap;
    buf.batch = batch_size;
  }
  float row[2][TGP_STRIDE] = {{0}}, box[2 * TGP_DIM];
  for (int k = 0; k < TGP_DIM; ++k) {
    row[0][k] = start[k];
    row[1][k] = goal[k];
    box[k] = lo[k];
    box[TGP_DIM + k] = hi[k];
  }
  const int two = 2;
  memset(p->status, 0, sizeof(TgPlanStatus));
  TG_CUDA(cudaMemcpyAsync(buf.nodes, row, sizeof(row), cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaMemcpyAsync(buf.count, &two, 4, cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaMemcpyAsync(p->box, box, sizeof(box), cudaMemcpyHostToDevice, st));
  TG_CUDA(cudaMemsetAsync(buf.ok, 0, 4 * (size_t)cap * words, st));
  TG_CUDA(cudaStreamSynchronize(st));
  const auto t0 = std::chrono::steady_clock::now();
  int drawn = 0, sweeps = 0, rc = TG_OK, n_old = 0;
  long long edges_total = 0;
  bool found = false;
  while (!found && drawn < max_samples) {
    // one more batch of samples joins the roadmap
    const int nb = drawn + batch_size <= max_samples ? batch_size : max_samples - drawn;
    k_fcit_samples<<<tg_div_up(nb, 256), 256, 0, st>>>(p->box, buf.nodes, halton_skip + (unsigned)drawn, nb, buf.arm, buf.base);
    rc = tg_check_configs(p->env, buf.arm, buf.base, nb, buf.valid, nullptr, st);
    if (rc) return rc;
    k_fcit_append<<<tg_div_up(nb, 256), 256, 0, st>>>(buf.valid, buf.arm, buf.base, nb, buf.nodes, buf.count, cap);
    drawn += nb;
    // every pair that involves a new node, closer than max_edge, becomes a candidate edge; validate them all at once
    TG_CUDA(cudaMemsetAsync(buf.n_edges, 0, 4, st));
    k_fcit_pairs<<<dim3(tg_div_up(cap - n_old, 256), cap), 256, 0, st>>>(buf.nodes, buf.count, n_old, max_edge, buf.e_arm0, buf.e_base0,
                                                                        buf.e_arm1, buf.e_base1, buf.pair, buf.n_edges, TGF_MAX_EDGES);
    k_plan_publish<<<1, 1, 0, st>>>(p->status_dev, buf.n_edges);
    TG_CUDA(cudaEventRecord(p->ev, st));
    TG_CUDA(cudaEventSynchronize(p->ev));
    int E = p->status->aux;  // a status word in mapped memory, not a copy
    if (E > TGF_MAX_EDGES) E = TGF_MAX_EDGES;
    if (E > 0) {
      rc = tg_check_edges(p->env, buf.e_arm0, buf.e_base0, buf.e_arm1, buf.e_base1, E, rake, resolution, buf.evalid, nullptr, st);
      if (rc) return rc;
      k_fcit_mark<<<tg_div_up(E, 256), 256, 0, st>>>(buf.evalid, buf.pair, E, buf.ok, words);
      edges_total += E;
    }
    k_plan_publish<<<1, 1, 0, st>>>(p->status_dev, buf.count);
    // cost-to-go over the validated graph
    k_fcit_init<<<tg_div_up(cap, 256), 256, 0, st>>>(buf.dist, buf.next, cap);
    for (int s = 0; s < max_rounds; ++s) {
      k_plan_clear<<<1, 1, 0, st>>>(p->status_dev, 1);
      for (int k = 0; k < 2; ++k)
        k_fcit_relax<<<tg_div_up(cap, 256), 256, 0, st>>>(p->status_dev, buf.nodes, buf.count, buf.ok, words, reverse_weight, max_edge, buf.dist,
                                                         buf.next);
      sweeps += 2;
      TG_CUDA(cudaEventRecord(p->ev, st));
      TG_CUDA(cudaEventSynchronize(p->ev));
      if (!p->status->changed) break;
    }
    n_old = p->status->aux;  // nodes now in the roadmap: the next batch only pairs its new nodes
    k_fcit_path<<<1, 1, 0, st>>>(p->status_dev, buf.nodes, buf.dist, buf.next, p->path, TGP_MAX_PATH);
    TG_CUDA(cudaEventRecord(p->ev, st));
    TG_CUDA(cudaEventSynchronize(p->ev));
    found = p->status->n_path >= 2;
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) break;
  }
  if (stats4) {
    stats4[0] = drawn;
    stats4[1] = n_old;
    stats4[2] = (double)edges_total;
    stats4[3] = sweeps;
  }
  TG_CUDA(cudaGetLastError());
  if (!found) return TG_OK;
  return tgp_fetch_path(p, path_out, max_path, n_path, st);
}
