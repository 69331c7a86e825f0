/*
 * cpu_checker.c — CPU restatement of the collision-verdict arithmetic (ORACLE / CPU BASELINE).
 *
 * TEST INFRASTRUCTURE: linked only by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs.  The product (tygrit_b200/) never loads it.
 *
 * What it restates: the VAMP-backed calls of TyGrit/planning/fetch/vamp_preview.py
 * (:159-163 validate_config, :306-338 check_whole_body_collisions / validate_whole_body_config,
 * :205-236 add_pointcloud).  VAMP itself (AdaCompNUS/vamp @ ab0bc39) is absent from the reference
 * tree, so this is a "port" of the published algorithm's semantics (sphere model vs point spheres,
 * fixed-resolution edge sampling), NOT VAMP's code: parity with VAMP is unpinned (see oracle/__init__.py).
 * FK follows TyGrit/kinematics/fetch/fk_numpy.py:65-155 and utils/transforms.py:116-127.
 *
 * Structure: dense uniform grid over the cloud box (counting sort), per-sphere scan of the cells that
 * overlap the sphere's box, early exit on the first hit, OpenMP over configurations; the fp32 verdict path is
 * coarse-to-fine (link by link along the chain, enclosing sphere before a link's spheres, largest first).
 *   tgo_check_*_f32   fp32 verdict path (what a CPU implementation would run; the timed baseline)
 *   tgo_clearance_f64 fp64 exact clearance (the oracle proper)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NL 15
#define NS 111

static const int LINK_PARENT[NL] = {-1, 0, 0, 2, 3, 2, 5, 6, 7, 8, 9, 10, 11, 12, 12};
/* 0 fixed, 1 prismatic z, 2 x, 3 y, 4 z */
static const int LINK_KIND[NL] = {0, 0, 1, 4, 3, 4, 3, 2, 3, 2, 3, 2, 0, 0, 0};
static const int LINK_Q[NL] = {-1, -1, 0, 8, 9, 1, 2, 3, 4, 5, 6, 7, -1, -1, -1};
static const double LINK_OFF[NL][3] = {
    {0, 0, 0},           {-0.086875, 0, 0.377425}, {-0.086875, 0, 0.37743}, {0.053125, 0, 0.603001},
    {0.14253, 0, 0.057999}, {0.119525, 0, 0.34858}, {0.117, 0, 0.06},       {0.219, 0, 0},
    {0.133, 0, 0},       {0.197, 0, 0},            {0.1245, 0, 0},          {0.1385, 0, 0},
    {0.16645, 0, 0},     {0, 0.065425, 0},         {0, -0.065425, 0}};

/* sphere table is handed over from oracle/fetch_model.py at load time (single source inside oracle/) */
static double SPH[NS][4];
static int SPH_LINK[NS];
static int model_set = 0;

/* derived at load time for the coarse-to-fine fp32 path: spheres of each link, largest first, and one
 * enclosing sphere per link (centre = mean of the link's sphere centres) */
static int LINK_NSPH[NL];
static int LINK_SPH[NL][32];
static double LINK_BOUND[NL][4];
/* optional self-collision term: sphere pairs of TyGrit/rl/train.py:78-107 SELF_COLLISION_PAIRS, grouped by link pair */
#define MAX_SELF_PAIRS 4096
static int SELF_I[MAX_SELF_PAIRS], SELF_J[MAX_SELF_PAIRS];
static int n_self = 0;

void tgo_set_model(const double* xyzr, const int64_t* link) {
  for (int i = 0; i < NS; ++i) {
    for (int k = 0; k < 4; ++k) SPH[i][k] = xyzr[4 * i + k];
    SPH_LINK[i] = (int)link[i];
  }
  for (int l = 0; l < NL; ++l) {
    int n = 0;
    double m[3] = {0, 0, 0};
    for (int i = 0; i < NS; ++i)
      if (SPH_LINK[i] == l) { LINK_SPH[l][n++] = i; for (int k = 0; k < 3; ++k) m[k] += SPH[i][k]; }
    LINK_NSPH[l] = n;
    for (int a = 1; a < n; ++a) { /* insertion sort, radius descending (stable) */
      const int v = LINK_SPH[l][a];
      int b = a - 1;
      while (b >= 0 && SPH[LINK_SPH[l][b]][3] < SPH[v][3]) { LINK_SPH[l][b + 1] = LINK_SPH[l][b]; --b; }
      LINK_SPH[l][b + 1] = v;
    }
    double rb = 0;
    for (int k = 0; k < 3; ++k) m[k] = n ? m[k] / n : 0;
    for (int a = 0; a < n; ++a) {
      const double* p = SPH[LINK_SPH[l][a]];
      const double d = sqrt((p[0] - m[0]) * (p[0] - m[0]) + (p[1] - m[1]) * (p[1] - m[1]) + (p[2] - m[2]) * (p[2] - m[2])) + p[3];
      if (d > rb) rb = d;
    }
    LINK_BOUND[l][0] = m[0]; LINK_BOUND[l][1] = m[1]; LINK_BOUND[l][2] = m[2]; LINK_BOUND[l][3] = rb * (1 + 1e-6) + 1e-6;
  }
  model_set = 1;
}

void tgo_set_self_pairs(const int64_t* pi, const int64_t* pj, int64_t n) {
  n_self = (int)(n < MAX_SELF_PAIRS ? n : MAX_SELF_PAIRS);
  for (int k = 0; k < n_self; ++k) { SELF_I[k] = (int)pi[k]; SELF_J[k] = (int)pj[k]; }
}

typedef struct {
  int64_t n;
  double lo[3];
  double cell;
  int dim[3];
  int64_t* start; /* dim0*dim1*dim2 + 1 */
  double* pd;     /* sorted points, fp64 */
  float* pf;      /* sorted points, fp32 */
  double r_pt;
} tgo_cloud;

void tgo_cloud_free(tgo_cloud* c) {
  if (!c) return;
  free(c->start); free(c->pd); free(c->pf); free(c);
}

tgo_cloud* tgo_cloud_build(const float* xyz, int64_t n, double r_pt, double cell) {
  tgo_cloud* c = (tgo_cloud*)calloc(1, sizeof(tgo_cloud));
  c->n = n; c->r_pt = r_pt; c->cell = cell;
  double hi[3] = {-1e300, -1e300, -1e300};
  c->lo[0] = c->lo[1] = c->lo[2] = 1e300;
  for (int64_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      double v = xyz[3 * i + a];
      if (v < c->lo[a]) c->lo[a] = v;
      if (v > hi[a]) hi[a] = v;
    }
  if (n == 0) { c->lo[0] = c->lo[1] = c->lo[2] = 0; hi[0] = hi[1] = hi[2] = 0; }
  int64_t nc = 1;
  for (int a = 0; a < 3; ++a) { c->dim[a] = (int)floor((hi[a] - c->lo[a]) / cell) + 1; nc *= c->dim[a]; }
  c->start = (int64_t*)calloc(nc + 1, sizeof(int64_t));
  int64_t* key = (int64_t*)malloc(sizeof(int64_t) * (n > 0 ? n : 1));
  for (int64_t i = 0; i < n; ++i) {
    int ix[3];
    for (int a = 0; a < 3; ++a) {
      ix[a] = (int)floor((xyz[3 * i + a] - c->lo[a]) / cell);
      if (ix[a] < 0) ix[a] = 0;
      if (ix[a] >= c->dim[a]) ix[a] = c->dim[a] - 1;
    }
    key[i] = ((int64_t)ix[2] * c->dim[1] + ix[1]) * c->dim[0] + ix[0];
    c->start[key[i] + 1]++;
  }
  for (int64_t k = 0; k < nc; ++k) c->start[k + 1] += c->start[k];
  int64_t* fill = (int64_t*)malloc(sizeof(int64_t) * (nc + 1));
  memcpy(fill, c->start, sizeof(int64_t) * (nc + 1));
  c->pd = (double*)malloc(sizeof(double) * 3 * (n > 0 ? n : 1));
  c->pf = (float*)malloc(sizeof(float) * 3 * (n > 0 ? n : 1));
  for (int64_t i = 0; i < n; ++i) {
    int64_t d = fill[key[i]]++;
    for (int a = 0; a < 3; ++a) { c->pd[3 * d + a] = xyz[3 * i + a]; c->pf[3 * d + a] = xyz[3 * i + a]; }
  }
  free(fill); free(key);
  return c;
}

#define DEFINE_IMPL(T, SFX, SIN, COS, SQRT, PTS)                                                                  \
  /* link frames in base_link: R[15][9], t[15][3] */                                                              \
  static void fk_one_##SFX(int i, const T* q, T R[NL][9], T t[NL][3]) {                                          \
    if (i == 0) {                                                                                                 \
      for (int k = 0; k < 9; ++k) R[0][k] = (k % 4 == 0) ? 1 : 0;                                                 \
      t[0][0] = t[0][1] = t[0][2] = 0;                                                                           \
      return;                                                                                                     \
    }                                                                                                             \
    {                                                                                                             \
      const int p = LINK_PARENT[i], kind = LINK_KIND[i];                                                          \
      T o[3] = {(T)LINK_OFF[i][0], (T)LINK_OFF[i][1], (T)LINK_OFF[i][2]};                                         \
      const T qi = LINK_Q[i] >= 0 ? q[LINK_Q[i]] : 0;                                                             \
      if (kind == 1) o[2] += qi;                                                                                  \
      for (int r = 0; r < 3; ++r) t[i][r] = R[p][3 * r] * o[0] + R[p][3 * r + 1] * o[1] + R[p][3 * r + 2] * o[2] + t[p][r]; \
      if (kind < 2) { memcpy(R[i], R[p], sizeof(T) * 9); return; }                                                \
      const T s = SIN(qi), c = COS(qi);                                                                           \
      const int a = kind == 2 ? 1 : (kind == 3 ? 2 : 0), b = kind == 2 ? 2 : (kind == 3 ? 0 : 1), k3 = 3 - a - b; \
      for (int r = 0; r < 3; ++r) {                                                                               \
        const T pa = R[p][3 * r + a], pb = R[p][3 * r + b];                                                       \
        R[i][3 * r + a] = c * pa + s * pb;                                                                        \
        R[i][3 * r + b] = c * pb - s * pa;                                                                        \
        R[i][3 * r + k3] = R[p][3 * r + k3];                                                                      \
      }                                                                                                           \
    }                                                                                                             \
  }                                                                                                               \
  static void fk_links_##SFX(const T* q, T R[NL][9], T t[NL][3]) {                                               \
    for (int i = 0; i < NL; ++i) fk_one_##SFX(i, q, R, t);                                                        \
  }                                                                                                               \
  /* world sphere centres (head joints given), out[111][3] */                                                     \
  static void centres_##SFX(const T* arm, const T* base, T hp, T ht, T out[NS][3]) {                             \
    T q[10], R[NL][9], t[NL][3];                                                                                  \
    for (int k = 0; k < 8; ++k) q[k] = arm[k];                                                                    \
    q[8] = hp; q[9] = ht;                                                                                         \
    fk_links_##SFX(q, R, t);                                                                                      \
    const T cb = COS(base[2]), sb = SIN(base[2]);                                                                 \
    for (int j = 0; j < NS; ++j) {                                                                                \
      const int l = SPH_LINK[j];                                                                                  \
      const T lx = (T)SPH[j][0], ly = (T)SPH[j][1], lz = (T)SPH[j][2];                                            \
      const T bx = R[l][0] * lx + R[l][1] * ly + R[l][2] * lz + t[l][0];                                          \
      const T by = R[l][3] * lx + R[l][4] * ly + R[l][5] * lz + t[l][1];                                          \
      const T bz = R[l][6] * lx + R[l][7] * ly + R[l][8] * lz + t[l][2];                                          \
      out[j][0] = cb * bx - sb * by + base[0];                                                                    \
      out[j][1] = sb * bx + cb * by + base[1];                                                                    \
      out[j][2] = bz;                                                                                             \
    }                                                                                                             \
  }                                                                                                               \
  /* signed clearance of one sphere vs the cloud, searching out to R + limit; returns limit if nothing closer */ \
  static T sphere_clear_##SFX(const tgo_cloud* c, const T* ctr, T R, T limit, int stop_at_hit) {                 \
    if (c->n == 0) return limit;                                                                                  \
    const T reach = R + limit;                                                                                    \
    int i0[3], i1[3];                                                                                             \
    for (int a = 0; a < 3; ++a) {                                                                                 \
      i0[a] = (int)floor(((double)ctr[a] - reach - c->lo[a]) / c->cell);                                          \
      i1[a] = (int)floor(((double)ctr[a] + reach - c->lo[a]) / c->cell);                                          \
      if (i0[a] < 0) i0[a] = 0;                                                                                   \
      if (i1[a] >= c->dim[a]) i1[a] = c->dim[a] - 1;                                                              \
      if (i0[a] > i1[a]) return limit;                                                                            \
    }                                                                                                             \
    T best2 = reach * reach;                                                                                      \
    const T R2 = R * R;                                                                                           \
    for (int z = i0[2]; z <= i1[2]; ++z)                                                                          \
      for (int y = i0[1]; y <= i1[1]; ++y) {                                                                      \
        const int64_t row = ((int64_t)z * c->dim[1] + y) * c->dim[0];                                             \
        const int64_t s = c->start[row + i0[0]], e = c->start[row + i1[0] + 1];                                   \
        for (int64_t p = s; p < e; ++p) {                                                                         \
          const T dx = (T)PTS[3 * p] - ctr[0], dy = (T)PTS[3 * p + 1] - ctr[1], dz = (T)PTS[3 * p + 2] - ctr[2]; \
          const T d2 = dx * dx + dy * dy + dz * dz;                                                               \
          if (stop_at_hit && d2 <= R2) return (T)-1;                                                              \
          if (d2 < best2) best2 = d2;                                                                             \
        }                                                                                                         \
      }                                                                                                           \
    return SQRT(best2) - R;                                                                                       \
  }

DEFINE_IMPL(double, f64, sin, cos, sqrt, c->pd)
DEFINE_IMPL(float, f32, sinf, cosf, sqrtf, c->pf)

/* ---- fp64 oracle: exact clearance per configuration, clipped at `cap` --------------------------- */
void tgo_clearance_f64(const tgo_cloud* c, const float* arm, const float* base, int64_t B, double cap, double* out) {
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t i = 0; i < B; ++i) {
    double a[8], b[3], ctr[NS][3];
    for (int k = 0; k < 8; ++k) a[k] = arm[8 * i + k];
    for (int k = 0; k < 3; ++k) b[k] = base[3 * i + k];
    centres_f64(a, b, 0.0, 0.0, ctr);
    double best = cap;
    for (int j = 0; j < NS; ++j) {
      const double R = SPH[j][3] + c->r_pt;
      const double v = sphere_clear_f64(c, ctr[j], R, best > 0 ? best : 0.0, 0);
      if (v < best) best = v;
    }
    out[i] = best;
  }
}

/* fp64 sphere centres (B x 111 x 3), head optional */
void tgo_centres_f64(const float* arm, const float* base, const float* head, int64_t B, double* out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < B; ++i) {
    double a[8], b[3];
    for (int k = 0; k < 8; ++k) a[k] = arm[8 * i + k];
    for (int k = 0; k < 3; ++k) b[k] = base[3 * i + k];
    centres_f64(a, b, head ? head[2 * i] : 0.0, head ? head[2 * i + 1] : 0.0, (double(*)[3])(out + (size_t)i * NS * 3));
  }
}

/* ---- optional self-collision term (TyGrit/rl/train.py:78-107 pairs): base_link frame, head = 0 ------ */
void tgo_self_clearance_f64(const float* arm, int64_t B, double* out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < B; ++i) {
    double a[8], b[3] = {0, 0, 0}, ctr[NS][3];
    for (int k = 0; k < 8; ++k) a[k] = arm[8 * i + k];
    centres_f64(a, b, 0.0, 0.0, ctr);
    double best = 1e300;
    for (int k = 0; k < n_self; ++k) {
      const int p = SELF_I[k], q = SELF_J[k];
      const double dx = ctr[p][0] - ctr[q][0], dy = ctr[p][1] - ctr[q][1], dz = ctr[p][2] - ctr[q][2];
      const double v = sqrt(dx * dx + dy * dy + dz * dz) - SPH[p][3] - SPH[q][3];
      if (v < best) best = v;
    }
    out[i] = best;
  }
}

/* ---- fp32 verdict path: the timed CPU baseline ---------------------------------------------------- */
/* Naive form (all 111 centres first, table order): kept as the second opinion of the tests. */
static int config_valid_naive_f32(const tgo_cloud* c, const float* arm, const float* base) {
  float ctr[NS][3];
  centres_f32(arm, base, 0.f, 0.f, ctr);
  for (int j = 0; j < NS; ++j) {
    const float R = (float)SPH[j][3] + (float)c->r_pt;
    if (sphere_clear_f32(c, ctr[j], R, 0.f, 1) < 0.f) return 0;
  }
  return 1;
}

/* Coarse-to-fine form, what a tuned scalar CPU checker does (and what VAMP's generated fkcc does per SIMD lane,
 * SURVEY.md 9.3): the link chain is advanced one link at a time and each link is tested as soon as its frame
 * exists - enclosing sphere first, then its spheres largest first - so a colliding base never pays for the arm FK.
 * Same fp32 sphere test, same verdicts as the naive form (the enclosing-sphere test is conservative). */
static int config_valid_f32(const tgo_cloud* c, const float* arm, const float* base, int self_on) {
  float q[10], R[NL][9], t[NL][3], W[NS][3];
  for (int k = 0; k < 8; ++k) q[k] = arm[k];
  q[8] = q[9] = 0.f;
  const float cb = cosf(base[2]), sb = sinf(base[2]);
  const float rpt = (float)c->r_pt;
  for (int l = 0; l < NL; ++l) {
    fk_one_f32(l, q, R, t);
    const int n = LINK_NSPH[l];
    if (n == 0) continue;
    int need = 1;
    if (n >= 3 && !self_on) { /* with the self term on, the world centres are needed later anyway */
      const float lx = (float)LINK_BOUND[l][0], ly = (float)LINK_BOUND[l][1], lz = (float)LINK_BOUND[l][2];
      const float bx = R[l][0] * lx + R[l][1] * ly + R[l][2] * lz + t[l][0];
      const float by = R[l][3] * lx + R[l][4] * ly + R[l][5] * lz + t[l][1];
      const float bz = R[l][6] * lx + R[l][7] * ly + R[l][8] * lz + t[l][2];
      const float w[3] = {cb * bx - sb * by + base[0], sb * bx + cb * by + base[1], bz};
      need = sphere_clear_f32(c, w, (float)LINK_BOUND[l][3] + rpt, 0.f, 1) < 0.f;
    }
    if (!need) continue;
    for (int a = 0; a < n; ++a) {
      const int j = LINK_SPH[l][a];
      const float lx = (float)SPH[j][0], ly = (float)SPH[j][1], lz = (float)SPH[j][2];
      const float bx = R[l][0] * lx + R[l][1] * ly + R[l][2] * lz + t[l][0];
      const float by = R[l][3] * lx + R[l][4] * ly + R[l][5] * lz + t[l][1];
      const float bz = R[l][6] * lx + R[l][7] * ly + R[l][8] * lz + t[l][2];
      W[j][0] = cb * bx - sb * by + base[0];
      W[j][1] = sb * bx + cb * by + base[1];
      W[j][2] = bz;
      if (sphere_clear_f32(c, W[j], (float)SPH[j][3] + rpt, 0.f, 1) < 0.f) return 0;
    }
  }
  if (self_on)
    for (int k = 0; k < n_self; ++k) {
      const int p = SELF_I[k], r = SELF_J[k];
      const float dx = W[p][0] - W[r][0], dy = W[p][1] - W[r][1], dz = W[p][2] - W[r][2];
      const float rr = (float)SPH[p][3] + (float)SPH[r][3];
      if (dx * dx + dy * dy + dz * dz <= rr * rr) return 0;
    }
  return 1;
}

/* mode bit 0: self-collision term on; bit 1: naive form (no staging) */
void tgo_check_configs_f32(const tgo_cloud* c, const float* arm, const float* base, int64_t B, uint8_t* valid, int threads, int mode) {
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
#endif
  if (mode & 2) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < B; ++i) valid[i] = (uint8_t)config_valid_naive_f32(c, arm + 8 * i, base + 3 * i);
    return;
  }
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t i = 0; i < B; ++i) valid[i] = (uint8_t)config_valid_f32(c, arm + 8 * i, base + 3 * i, mode & 1);
}

/* edges: n = max(ceil(|d| * resolution / rake), 1) in fp64 (order x,y,theta,q0..q7); samples k/(rake n) */
static int64_t edge_total(const float* a0, const float* b0, const float* a1, const float* b1, int rake, double resolution) {
  double s2 = 0;
  for (int k = 0; k < 3; ++k) { const double d = (double)b1[k] - (double)b0[k]; s2 = s2 + d * d; }
  for (int k = 0; k < 8; ++k) { const double d = (double)a1[k] - (double)a0[k]; s2 = s2 + d * d; }
  double n = ceil(sqrt(s2) * (resolution / rake));
  if (n < 1) n = 1;
  return (int64_t)n * rake;
}

int64_t tgo_check_edges_f32(const tgo_cloud* c, const float* arm0, const float* base0, const float* arm1, const float* base1,
                            int64_t E, int rake, double resolution, uint8_t* valid, int threads, int mode) {
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
#endif
  int64_t checks = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : checks)
  for (int64_t e = 0; e < E; ++e) {
    const int64_t total = edge_total(arm0 + 8 * e, base0 + 3 * e, arm1 + 8 * e, base1 + 3 * e, rake, resolution);
    checks += total;
    int ok = 1;
    /* VAMP-style order: stride through the segment so coarse coverage comes first */
    const int64_t n = total / rake;
    for (int64_t blk = 0; blk < n && ok; ++blk)
      for (int r = 0; r < rake && ok; ++r) {
        const int64_t k = (int64_t)r * n + blk + 1;
        const float t = (float)((double)k / (double)total);
        float a[8], b[3];
        for (int j = 0; j < 8; ++j) a[j] = arm0[8 * e + j] + t * (arm1[8 * e + j] - arm0[8 * e + j]);
        for (int j = 0; j < 3; ++j) b[j] = base0[3 * e + j] + t * (base1[3 * e + j] - base0[3 * e + j]);
        ok = config_valid_f32(c, a, b, mode & 1);
      }
    valid[e] = (uint8_t)ok;
  }
  return checks;
}

/* fp64 min clearance over the samples of each edge (oracle for edge verdicts) */
void tgo_edge_clearance_f64(const tgo_cloud* c, const float* arm0, const float* base0, const float* arm1, const float* base1,
                            int64_t E, int rake, double resolution, double cap, double* out, int64_t* nsamp) {
#pragma omp parallel for schedule(dynamic, 8)
  for (int64_t e = 0; e < E; ++e) {
    const int64_t total = edge_total(arm0 + 8 * e, base0 + 3 * e, arm1 + 8 * e, base1 + 3 * e, rake, resolution);
    if (nsamp) nsamp[e] = total;
    double best = cap;
    for (int64_t k = 1; k <= total; ++k) {
      const double t = (double)k / (double)total;
      double a[8], b[3], ctr[NS][3];
      for (int j = 0; j < 8; ++j) a[j] = (double)arm0[8 * e + j] + t * ((double)arm1[8 * e + j] - (double)arm0[8 * e + j]);
      for (int j = 0; j < 3; ++j) b[j] = (double)base0[3 * e + j] + t * ((double)base1[3 * e + j] - (double)base0[3 * e + j]);
      centres_f64(a, b, 0.0, 0.0, ctr);
      for (int j = 0; j < NS; ++j) {
        const double v = sphere_clear_f64(c, ctr[j], SPH[j][3] + c->r_pt, best > 0 ? best : 0.0, 0);
        if (v < best) best = v;
      }
    }
    out[e] = best;
  }
}

/* OpenMP thread count for every later call (0 = leave as is); torchrun exports OMP_NUM_THREADS=1, which the
 * CPU baseline must not inherit silently */
void tgo_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
int tgo_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
int tgo_model_ready(void) { return model_set; }
