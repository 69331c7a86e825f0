// query.cuh — device functions of the fused FK + sphere-query path (K2 + K3).
//
// Phase 1 (tg_stage_a / tg_stage_b): one lane = one configuration.  The planner's robot model has
// the head fixed, so 55 of the 111 spheres (base, torso, head) only need the SE(2) base pose and
// the torso lift; the remaining 56 ride the 7-joint arm chain.  Every sphere centre is produced
// directly in distance-grid coordinates (voxels) and costs ONE byte load from the L2-resident grid,
// giving a three-way answer per sphere: definitely free / definitely colliding / ambiguous.
// The kernel is bound by the number of distinct L1/L2 sectors a warp's 32 random probes touch
// (about one per clock per SM), so the spheres of every frame are ordered largest first and probed
// in small groups: a lane whose configuration is already refuted issues no further loads.
// Phase 2 (tg_warp_sphere_hits): exact test of an ambiguous sphere against the sorted cell rows,
// one warp per sphere — lanes take the (z, y) cell rows, then scan the candidate points of all
// rows together with coalesced 16-byte loads and leave through __any_sync on the first hit.
#pragma once
#include "common.cuh"
#include "fk.cuh"

#ifdef TG_STATS
__device__ unsigned long long tg_stats[16];
#define TG_STAT(i, v) do { if ((threadIdx.x & 31) == 0) atomicAdd(&tg_stats[i], (unsigned long long)(v)); } while (0)
#define TG_STAT_LANE(i, v) atomicAdd(&tg_stats[i], (unsigned long long)(v))  // the caller picks the lane
#else
#define TG_STAT(i, v) do { } while (0)
#define TG_STAT_LANE(i, v) do { } while (0)
#endif

// Distance code of the voxel containing (gx, gy, gz) (voxel units).  Adding the 2^23 bias rounds to the
// nearest voxel in the mantissa; subtracting the bit pattern of 2^23 as an unsigned and taking the
// unsigned minimum with n - 1 clamps both sides at once — one DPX add-min per axis — because values
// below the grid wrap to huge unsigned numbers.  They land on the far boundary voxel instead of the near
// one, which is equivalent: every boundary voxel is at least W voxels from the cloud (tg_plan_grids pads
// the box by W + 3), so both read "free".
__device__ __forceinline__ int tg_probe(const TgGrid& g, float gx, float gy, float gz) {
  TG_STAT_LANE(10, 1);  // executed probes (one per lane)
#ifndef TG_LINEAR_PROBE
  // texture path: point sampling floors (g - m + 0.5) = the nearest voxel, clamp addressing keeps probes outside the
  // grid on a boundary voxel (which reads "free": the box is padded by W + 3 voxels), block-linear array layout
  // (PTX rather than tex3D<unsigned char>: one component requested, no conversion - 1 % fewer instructions issued)
  unsigned c;
  asm("{\n\t.reg .u32 t1, t2, t3;\n\ttex.3d.v4.u32.f32 {%0, t1, t2, t3}, [%1, {%2, %3, %4, %4}];\n\t}"
      : "=r"(c)
      : "l"(g.df_tex), "f"(gx + g.tex_bias[0]), "f"(gy + g.tex_bias[1]), "f"(gz + g.tex_bias[2]));
  return (int)c;
#endif
  const unsigned ux = __viaddmin_u32(__float_as_uint(gx + g.bias[0]), 0u - TG_MAGIC_BITS, (unsigned)(g.nx - 1));
  const unsigned uy = __viaddmin_u32(__float_as_uint(gy + g.bias[1]), 0u - TG_MAGIC_BITS, (unsigned)(g.ny - 1));
  const unsigned uz = __viaddmin_u32(__float_as_uint(gz + g.bias[2]), 0u - TG_MAGIC_BITS, (unsigned)(g.nz - 1));
  return (int)__ldg(g.df + (ux + (unsigned)g.nx * uy + (unsigned)g.nxy * uz));
}

// Predicated probe: lanes with need == false issue no request and read `skipped`.  Written as predicated PTX inside
// straight-line code so the texture handle stays in a uniform register (a tex3D under a divergent branch makes the
// compiler wrap every fetch in a handle-uniformity loop).
__device__ __forceinline__ int tg_probe_if(const TgGrid& g, bool need, float gx, float gy, float gz, int skipped) {
#ifdef TG_STATS
  atomicAdd(&tg_stats[need ? 10 : 11], 1ull);  // executed / skipped (derived bound) probes
#endif
#ifndef TG_LINEAR_PROBE
  unsigned c;
  asm("{\n\t.reg .pred p;\n\t.reg .u32 t1, t2, t3;\n\tsetp.ne.s32 p, %5, 0;\n\tmov.u32 %0, %6;\n\t"
      "@p tex.3d.v4.u32.f32 {%0, t1, t2, t3}, [%1, {%2, %3, %4, %4}];\n\t}"
      : "=&r"(c)
      : "l"(g.df_tex), "f"(gx + g.tex_bias[0]), "f"(gy + g.tex_bias[1]), "f"(gz + g.tex_bias[2]), "r"((int)need), "r"(skipped));
  return (int)c;
#else
  return need ? tg_probe(g, gx, gy, gz) : skipped;
#endif
}

// Coarse far-field grid in SHARED memory (phase 1 of the configuration kernel).  One byte per block of cg_r^3 voxels =
// a lower bound on the distance from any point of the block to the cloud (K1, k_edt_z_coarse).  A sphere whose centre
// reads a code >= thr.cfree_ge[k] is proven free without touching the L2-resident fine grid: the look-up is ~12 ALU
// instructions and one LDS, while a fine probe costs one L1-miss request - the resource that bounds this kernel.
struct TgCoarse {
  const uint8_t* s;  // shared-memory copy, [gnz][gny][gnx]
  int nx1, ny1, nz1; // gn - 1
  int nx, nxy;
  float inv_r, b0, b1, b2;  // block index = floor((g + tex_bias) * inv_r)
};
#define TG_FLOOR_MAGIC 12582912.0f      // 1.5 * 2^23: x + this, rounded DOWN, holds floor(x) in its mantissa for |x| < 2^22
#define TG_FLOOR_MAGIC_BITS 0x4B400000
__device__ __forceinline__ TgCoarse tg_coarse_init(const TgGrid& g, const uint8_t* smem) {
  TgCoarse c;
  c.s = smem;
  c.nx1 = g.gnx - 1; c.ny1 = g.gny - 1; c.nz1 = g.gnz - 1;
  c.nx = g.gnx; c.nxy = g.gnx * g.gny;
  c.inv_r = g.cg_inv_r;
  c.b0 = g.tex_bias[0] * g.cg_inv_r + TG_FLOOR_MAGIC; c.b1 = g.tex_bias[1] * g.cg_inv_r + TG_FLOOR_MAGIC;
  c.b2 = g.tex_bias[2] * g.cg_inv_r + TG_FLOOR_MAGIC;
  return c;
}
__device__ __forceinline__ int tg_coarse_code(const TgCoarse& c, float gx, float gy, float gz) {
  // Block index per axis = floor(g * inv_r + b): ONE round-down FMA with the 1.5 * 2^23 bias folded into b leaves the
  // integer in the mantissa (no F2I: the conversion unit runs at quarter rate and this kernel is issue-bound), and ONE
  // DPX add-min-relu subtracts the bias pattern and clamps to [0, n - 1].  Points outside the grid read the nearest
  // boundary block, whose bound holds for them too (projecting onto the box does not increase the distance to a point
  // inside it).  |g * inv_r| stays far below 2^22.
  const int ix = __viaddmin_s32_relu(__float_as_int(__fmaf_rd(gx, c.inv_r, c.b0)), -TG_FLOOR_MAGIC_BITS, c.nx1);
  const int iy = __viaddmin_s32_relu(__float_as_int(__fmaf_rd(gy, c.inv_r, c.b1)), -TG_FLOOR_MAGIC_BITS, c.ny1);
  const int iz = __viaddmin_s32_relu(__float_as_int(__fmaf_rd(gz, c.inv_r, c.b2)), -TG_FLOOR_MAGIC_BITS, c.nz1);
  TG_STAT_LANE(13, 1);
  return (int)c.s[iz * c.nxy + iy * c.nx + ix];
}

#ifndef TG_COARSE_MEMBERS
#define TG_COARSE_MEMBERS 0  // 1: members of a cull group whose enclosing sphere was not proven free also try the coarse grid
#endif

struct TgClass {
  bool collide;
  uint32_t amb[4];
};

__device__ __forceinline__ void tg_class_clear(TgClass& c) {
  c.collide = false;
  c.amb[0] = c.amb[1] = c.amb[2] = c.amb[3] = 0u;
}

#define TG_CLASSIFY_ONE(k, code)                                          \
  do {                                                                    \
    const int _c = (code);                                                \
    out.collide |= _c <= thr.coll_le[k];                                  \
    if (_c < thr.free_ge[k]) out.amb[(k) >> 5] |= 1u << ((k) & 31);       \
  } while (0)

// base pose in voxel units: G = vpm * Rz(theta), gt = vpm * (x, y)
struct TgPose {
  float cs, sn, gx0, gy0;
};
__device__ __forceinline__ TgPose tg_pose(const TgGrid& g, float bx, float by, float bth) {
  float s, c;
  tg_sincos(bth, &s, &c);
  TgPose p;
  p.cs = g.vpm * c;
  p.sn = g.vpm * s;
  p.gx0 = g.vpm * bx;
  p.gy0 = g.vpm * by;
  return p;
}

// Derived bounds (phase 1 of the configuration kernel; off with -DTG_NO_DERIVED_BOUNDS): a member k of a cull group is
// not probed at all when the code already read at the centre of the group's enclosing sphere (its "source",
// fetch_model.cuh tgm_k_src) is >= thr.src_ge[k], i.e. proves the ball of radius |c_k - c_src| + r_k + r_pt around the
// source free - a per-sphere version of the all-or-nothing group test.  A skipped sphere reads TG_CODE_SKIPPED: free,
// never ambiguous.  The edge kernel does not use it (measured 8 % slower there: more predicates, same loads).
#define TG_CODE_SKIPPED (1 << 20)
template <int K, bool DERIVE, bool C>
__device__ __forceinline__ int tg_probe_k(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, int bcode, float gx, float gy, float gz) {
  bool need = true;
#ifndef TG_NO_DERIVED_BOUNDS
  if constexpr (DERIVE && tgm_k_src(K) >= 0) need = bcode < thr.src_ge[K];
#endif
  if constexpr (C) {
    need = need && tg_coarse_code(cg, gx, gy, gz) < thr.cfree_ge[K];
    return tg_probe_if(g, need, gx, gy, gz, TG_CODE_SKIPPED);
  } else {
#ifndef TG_NO_DERIVED_BOUNDS
    if constexpr (DERIVE && tgm_k_src(K) >= 0) return tg_probe_if(g, need, gx, gy, gz, TG_CODE_SKIPPED);
#endif
    return tg_probe(g, gx, gy, gz);
  }
}

// spheres [K0, K1) of a frame whose rotation is the base yaw only:  p_base = l + (tx, 0, tz).
// All loads of the group are issued before the first classification (memory-level parallelism).
// DERIVE: bcode is the code read at the centre of the group's enclosing sphere.
template <int K, int K1, int K0, bool DERIVE, bool C>
__device__ __forceinline__ void tg_probe_yaw(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, float gtx, float gty,
                                             float gtz, int bcode, int (&code)[K1 - K0]) {
  if constexpr (K < K1) {
    const float4 l = TGM_K_SPH[K];
    const float gx = fmaf(p.cs, l.x, fmaf(-p.sn, l.y, gtx));
    const float gy = fmaf(p.sn, l.x, fmaf(p.cs, l.y, gty));
    const float gz = fmaf(g.vpm, l.z, gtz);
    code[K - K0] = tg_probe_k<K, DERIVE, C>(g, thr, cg, bcode, gx, gy, gz);
    tg_probe_yaw<K + 1, K1, K0, DERIVE, C>(g, thr, cg, p, gtx, gty, gtz, bcode, code);
  }
}
template <int K0, int K1, bool DERIVE, bool C>
__device__ __forceinline__ void tg_group_yaw(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, float gtx, float gty,
                                             float gtz, TgClass& out, int bcode = 0) {
  int code[K1 - K0];
  tg_probe_yaw<K0, K1, K0, DERIVE, C>(g, thr, cg, p, gtx, gty, gtz, bcode, code);
#pragma unroll
  for (int k = K0; k < K1; ++k) TG_CLASSIFY_ONE(k, code[k - K0]);
}

// The same for a lane that may be switched off (`en` = false: no look-up, no probe, nothing classified).  Branch-free, so
// that the calls for several configurations of one thread (stage A1 of the configuration kernel runs TGC_ILP of them per
// lane) are one basic block and their probes are in flight together.
template <int K0, int K1, bool C>
__device__ __forceinline__ void tg_group_yaw_en(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, bool en, TgClass& out) {
  int code[K1 - K0];
#pragma unroll
  for (int k = K0; k < K1; ++k) {
    const float4 l = TGM_K_SPH[k];
    const float gx = fmaf(p.cs, l.x, fmaf(-p.sn, l.y, p.gx0));
    const float gy = fmaf(p.sn, l.x, fmaf(p.cs, l.y, p.gy0));
    const float gz = g.vpm * l.z;
    bool need = en;
    if constexpr (C) need = need && tg_coarse_code(cg, gx, gy, gz) < thr.cfree_ge[k];
    code[k - K0] = tg_probe_if(g, need, gx, gy, gz, TG_CODE_SKIPPED);  // a skipped sphere reads "free, not ambiguous"
  }
#pragma unroll
  for (int k = K0; k < K1; ++k) TG_CLASSIFY_ONE(k, code[k - K0]);
}

// voxel-space affine map of an arm frame F (base_link frame) under base pose p
struct TgVox {
  float G[9], gt[3];
};
__device__ __forceinline__ void tg_vox_frame(const TgGrid& g, const TgPose& p, const TgFrame& F, TgVox& V) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    V.G[c] = fmaf(p.cs, F.r[c], -p.sn * F.r[3 + c]);
    V.G[3 + c] = fmaf(p.sn, F.r[c], p.cs * F.r[3 + c]);
    V.G[6 + c] = g.vpm * F.r[6 + c];
  }
  V.gt[0] = fmaf(p.cs, F.t[0], fmaf(-p.sn, F.t[1], p.gx0));
  V.gt[1] = fmaf(p.sn, F.t[0], fmaf(p.cs, F.t[1], p.gy0));
  V.gt[2] = g.vpm * F.t[2];
}

template <int K, int K1, int K0, bool DERIVE, bool C>
__device__ __forceinline__ void tg_probe_arm(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgVox& V, int bcode,
                                             int (&code)[K1 - K0]) {
  if constexpr (K < K1) {
    const float4 l = TGM_K_SPH[K];
    const float gx = fmaf(V.G[0], l.x, fmaf(V.G[1], l.y, fmaf(V.G[2], l.z, V.gt[0])));
    const float gy = fmaf(V.G[3], l.x, fmaf(V.G[4], l.y, fmaf(V.G[5], l.z, V.gt[1])));
    const float gz = fmaf(V.G[6], l.x, fmaf(V.G[7], l.y, fmaf(V.G[8], l.z, V.gt[2])));
    code[K - K0] = tg_probe_k<K, DERIVE, C>(g, thr, cg, bcode, gx, gy, gz);
    tg_probe_arm<K + 1, K1, K0, DERIVE, C>(g, thr, cg, V, bcode, code);
  }
}
template <int K0, int K1, bool DERIVE, bool C>
__device__ __forceinline__ void tg_group_arm(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgVox& V, int bcode, TgClass& out) {
  int code[K1 - K0];
  tg_probe_arm<K0, K1, K0, DERIVE, C>(g, thr, cg, V, bcode, code);
#pragma unroll
  for (int k = K0; k < K1; ++k) TG_CLASSIFY_ONE(k, code[k - K0]);
}

// enclosing-sphere probes (cull groups, fetch_model.cuh TGM_BOUND): true = the whole group is free
template <int B, bool C>
__device__ __forceinline__ int tg_bound_code_yaw(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, float gtx, float gty,
                                                 float gtz) {
  const float4 l = TGM_BOUND[B];
  const float gx = fmaf(p.cs, l.x, fmaf(-p.sn, l.y, gtx)), gy = fmaf(p.sn, l.x, fmaf(p.cs, l.y, gty)), gz = fmaf(g.vpm, l.z, gtz);
  if constexpr (C) return tg_probe_if(g, tg_coarse_code(cg, gx, gy, gz) < thr.cbound_ge[B], gx, gy, gz, TG_CODE_SKIPPED);
  return tg_probe(g, gx, gy, gz);
}
// Stage A: the 20 spheres that only depend on the base pose (base_link + torso_fixed), largest first.
// Most colliding configurations are refuted by the first group of three.
// Stage A is split in two: the three big base spheres (they refute 65 % of the bench's configurations) and the other 17.
// The configuration kernel compacts its survivors between the two parts (TG_EARLY_COMPACT), so the 17 smaller spheres
// already run with full warps; the edge kernel runs both parts back to back.
template <bool C>
__device__ __forceinline__ void tg_stage_a1(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, TgClass& out) {
  // the largest base sphere alone refutes 55 % of the bench's configurations, the three big ones 65 %
  tg_group_yaw<0, 1, false, C>(g, thr, cg, p, p.gx0, p.gy0, 0.f, out);
  if (out.collide) return;
  tg_group_yaw<1, 3, false, C>(g, thr, cg, p, p.gx0, p.gy0, 0.f, out);
}
template <bool C>
__device__ __forceinline__ void tg_stage_a2(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, TgClass& out) {
  // (enclosing-sphere probes for the torso column and the bumper arc were measured: they rarely prove anything inside
  // a 0.32 m window and cost 1 % more than they save)
#ifndef TG_A2_SPLIT  // all 17 probes in flight together: one latency step less for a few more probes (phase 1 1.81 -> 1.77 ms; -DTG_A2_SPLIT: the two groups with a lane-level exit between them)
  tg_group_yaw<3, 20, false, C>(g, thr, cg, p, p.gx0, p.gy0, 0.f, out);
#else
  tg_group_yaw<3, 9, false, C>(g, thr, cg, p, p.gx0, p.gy0, 0.f, out);
  if (out.collide) return;
  tg_group_yaw<9, 20, false, C>(g, thr, cg, p, p.gx0, p.gy0, 0.f, out);
#endif
}
template <bool C>
__device__ __forceinline__ void tg_stage_a(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, TgClass& out) {
  tg_stage_a1<C>(g, thr, cg, p, out);
  if (out.collide) return;
  tg_stage_a2<C>(g, thr, cg, p, out);
}

template <int B, bool C>
__device__ __forceinline__ int tg_bound_code_arm(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgVox& V) {
  const float4 l = TGM_BOUND[B];
  const float gx = fmaf(V.G[0], l.x, fmaf(V.G[1], l.y, fmaf(V.G[2], l.z, V.gt[0])));
  const float gy = fmaf(V.G[3], l.x, fmaf(V.G[4], l.y, fmaf(V.G[5], l.z, V.gt[1])));
  const float gz = fmaf(V.G[6], l.x, fmaf(V.G[7], l.y, fmaf(V.G[8], l.z, V.gt[2])));
  if constexpr (C) return tg_probe_if(g, tg_coarse_code(cg, gx, gy, gz) < thr.cbound_ge[B], gx, gy, gz, TG_CODE_SKIPPED);
  return tg_probe(g, gx, gy, gz);
}

// The arm chain in two passes (-DTG_ARM_TWO_PASS).  The chain as written in tg_stage_b is up to 14 dependent probe
// latencies (per link: the enclosing-sphere probe, then the members if it did not prove them free).  Here pass 1 runs the
// FK chain once and issues the seven enclosing-sphere probes without looking at them (they are in flight together), and
// pass 2 - only for lanes with at least one group to look into - runs the chain again up to the last such link and probes
// the members there.  Same probes, same classification; the second FK pass costs ~400 instructions per warp that needs it.
// MEASURED VARIANT, not the default: parity-green and 20 % SLOWER (phase 1 1.777 -> 2.138 ms per 32 Mi configurations, free-heavy
// batch 4.25 -> 4.92 ms, C2 0.622 -> 0.667 ms; profiles/r03n_arm_two_pass.txt).  Unlike stage A1, stage B is not waiting for its
// probes: what it lacks is issue slots, and the second FK pass takes more of them.
#define TG_ARM_CHILD(dst, src, NAME, kind, qi) \
  tg_frame_child(src, TGM_OFF_##NAME##_X, TGM_OFF_##NAME##_Y, TGM_OFF_##NAME##_Z, kind, q[qi], dst)
template <bool DERIVE, bool C>
__device__ __forceinline__ void tg_arm_two_pass(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, const float* q,
                                                float tz, TgClass& out) {
  TgFrame T, F, H;
  TgVox V;
  tg_frame_identity(T);
  T.t[0] = TGM_OFF_TORSO_X;
  T.t[2] = tz;
  int bc1, bc2, bc3, bc4, bc5, bc6, bc7;
  TG_ARM_CHILD(F, T, SHOULDER_PAN, 4, 1);  tg_vox_frame(g, p, F, V); bc1 = tg_bound_code_arm<3, C>(g, thr, cg, V);
  TG_ARM_CHILD(H, F, SHOULDER_LIFT, 3, 2); tg_vox_frame(g, p, H, V); bc2 = tg_bound_code_arm<4, C>(g, thr, cg, V);
  TG_ARM_CHILD(F, H, UPPERARM_ROLL, 2, 3); tg_vox_frame(g, p, F, V); bc3 = tg_bound_code_arm<5, C>(g, thr, cg, V);
  TG_ARM_CHILD(H, F, ELBOW_FLEX, 3, 4);    tg_vox_frame(g, p, H, V); bc4 = tg_bound_code_arm<6, C>(g, thr, cg, V);
  TG_ARM_CHILD(F, H, FOREARM_ROLL, 2, 5);  tg_vox_frame(g, p, F, V); bc5 = tg_bound_code_arm<7, C>(g, thr, cg, V);
  TG_ARM_CHILD(H, F, WRIST_FLEX, 3, 6);    tg_vox_frame(g, p, H, V); bc6 = tg_bound_code_arm<8, C>(g, thr, cg, V);
  TG_ARM_CHILD(F, H, WRIST_ROLL, 2, 7);    tg_vox_frame(g, p, F, V); bc7 = tg_bound_code_arm<9, C>(g, thr, cg, V);
  unsigned need = (bc1 < thr.bound_ge[3] ? 1u : 0u) | (bc2 < thr.bound_ge[4] ? 2u : 0u) | (bc3 < thr.bound_ge[5] ? 4u : 0u) |
                  (bc4 < thr.bound_ge[6] ? 8u : 0u) | (bc5 < thr.bound_ge[7] ? 16u : 0u) | (bc6 < thr.bound_ge[8] ? 32u : 0u);
  if (bc7 < thr.bound_ge[9]) {  // the last link's frame is still here
    tg_group_arm<93, 99, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc7, out);
    if (out.collide) return;
    tg_group_arm<99, 111, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc7, out);
    if (out.collide) return;
  }
  if (need == 0u) return;
  TG_ARM_CHILD(F, T, SHOULDER_PAN, 4, 1);
  if (need & 1u) {
    tg_vox_frame(g, p, F, V);
    tg_group_arm<55, 59, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc1, out);
    if (out.collide) return;
  }
  if ((need >> 1) == 0u) return;
  TG_ARM_CHILD(H, F, SHOULDER_LIFT, 3, 2);
  if (need & 2u) {
    tg_vox_frame(g, p, H, V);
    tg_group_arm<59, 66, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc2, out);
    if (out.collide) return;
  }
  if ((need >> 2) == 0u) return;
  TG_ARM_CHILD(F, H, UPPERARM_ROLL, 2, 3);
  if (need & 4u) {
    tg_vox_frame(g, p, F, V);
    tg_group_arm<66, 74, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc3, out);
    if (out.collide) return;
  }
  if ((need >> 3) == 0u) return;
  TG_ARM_CHILD(H, F, ELBOW_FLEX, 3, 4);
  if (need & 8u) {
    tg_vox_frame(g, p, H, V);
    tg_group_arm<74, 80, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc4, out);
    if (out.collide) return;
  }
  if ((need >> 4) == 0u) return;
  TG_ARM_CHILD(F, H, FOREARM_ROLL, 2, 5);
  if (need & 16u) {
    tg_vox_frame(g, p, F, V);
    tg_group_arm<80, 87, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc5, out);
    if (out.collide) return;
  }
  if ((need >> 5) == 0u) return;
  TG_ARM_CHILD(H, F, WRIST_FLEX, 3, 6);
  tg_vox_frame(g, p, H, V);
  tg_group_arm<87, 93, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc6, out);
}

// Stage B: torso + head (35 spheres, base yaw + torso lift) and the 7-joint arm chain (56 spheres).
// The small spheres are grouped (head, ears, one group per arm link); one probe at the centre of a
// group's enclosing sphere that proves it free skips the whole group, which is the common case for a
// free configuration away from the surfaces.
template <bool DERIVE, bool C>
__device__ __forceinline__ void tg_stage_b(const TgGrid& g, const TgThr& thr, const TgCoarse& cg, const TgPose& p, const float* q, TgClass& out) {
  const float tz = TGM_OFF_TORSO_Z + q[0];
  {
    const float gtx1 = fmaf(p.cs, TGM_OFF_TORSO_X, p.gx0), gty1 = fmaf(p.sn, TGM_OFF_TORSO_X, p.gy0), gtz1 = g.vpm * tz;
    const int b0 = tg_bound_code_yaw<0, C>(g, thr, cg, p, gtx1, gty1, gtz1);
    const int b1 = tg_bound_code_yaw<1, C>(g, thr, cg, p, gtx1, gty1, gtz1);
    const int b2 = tg_bound_code_yaw<2, C>(g, thr, cg, p, gtx1, gty1, gtz1);
    tg_group_yaw<20, 28, false, C>(g, thr, cg, p, gtx1, gty1, gtz1, out);
    if (out.collide) return;
    if (b0 < thr.bound_ge[0]) {
      tg_group_yaw<28, 33, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, p, gtx1, gty1, gtz1, out, b0);
      if (out.collide) return;
    }
    if (b1 < thr.bound_ge[1]) {
      tg_group_yaw<33, 44, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, p, gtx1, gty1, gtz1, out, b1);
      if (out.collide) return;
    }
    if (b2 < thr.bound_ge[2]) {
      tg_group_yaw<44, 55, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, p, gtx1, gty1, gtz1, out, b2);
      if (out.collide) return;
    }
  }
#ifdef TG_ARM_TWO_PASS
  tg_arm_two_pass<DERIVE, C>(g, thr, cg, p, q, tz, out);
}
#else
  TgFrame T, F, H;
  TgVox V;
  tg_frame_identity(T);
  T.t[0] = TGM_OFF_TORSO_X;
  T.t[2] = tz;
  tg_frame_child(T, TGM_OFF_SHOULDER_PAN_X, TGM_OFF_SHOULDER_PAN_Y, TGM_OFF_SHOULDER_PAN_Z, 4, q[1], F);
  tg_vox_frame(g, p, F, V);
  if (const int bc = tg_bound_code_arm<3, C>(g, thr, cg, V); bc < thr.bound_ge[3]) {
    tg_group_arm<55, 59, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(F, TGM_OFF_SHOULDER_LIFT_X, TGM_OFF_SHOULDER_LIFT_Y, TGM_OFF_SHOULDER_LIFT_Z, 3, q[2], H);
  tg_vox_frame(g, p, H, V);
  if (const int bc = tg_bound_code_arm<4, C>(g, thr, cg, V); bc < thr.bound_ge[4]) {
    tg_group_arm<59, 66, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(H, TGM_OFF_UPPERARM_ROLL_X, TGM_OFF_UPPERARM_ROLL_Y, TGM_OFF_UPPERARM_ROLL_Z, 2, q[3], F);
  tg_vox_frame(g, p, F, V);
  if (const int bc = tg_bound_code_arm<5, C>(g, thr, cg, V); bc < thr.bound_ge[5]) {
    tg_group_arm<66, 74, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(F, TGM_OFF_ELBOW_FLEX_X, TGM_OFF_ELBOW_FLEX_Y, TGM_OFF_ELBOW_FLEX_Z, 3, q[4], H);
  tg_vox_frame(g, p, H, V);
  if (const int bc = tg_bound_code_arm<6, C>(g, thr, cg, V); bc < thr.bound_ge[6]) {
    tg_group_arm<74, 80, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(H, TGM_OFF_FOREARM_ROLL_X, TGM_OFF_FOREARM_ROLL_Y, TGM_OFF_FOREARM_ROLL_Z, 2, q[5], F);
  tg_vox_frame(g, p, F, V);
  if (const int bc = tg_bound_code_arm<7, C>(g, thr, cg, V); bc < thr.bound_ge[7]) {
    tg_group_arm<80, 87, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(F, TGM_OFF_WRIST_FLEX_X, TGM_OFF_WRIST_FLEX_Y, TGM_OFF_WRIST_FLEX_Z, 3, q[6], H);
  tg_vox_frame(g, p, H, V);
  if (const int bc = tg_bound_code_arm<8, C>(g, thr, cg, V); bc < thr.bound_ge[8]) {
    tg_group_arm<87, 93, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
  }
  tg_frame_child(H, TGM_OFF_WRIST_ROLL_X, TGM_OFF_WRIST_ROLL_Y, TGM_OFF_WRIST_ROLL_Z, 2, q[7], F);
  tg_vox_frame(g, p, F, V);
  if (const int bc = tg_bound_code_arm<9, C>(g, thr, cg, V); bc < thr.bound_ge[9]) {
    tg_group_arm<93, 99, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
    if (out.collide) return;
    tg_group_arm<99, 111, DERIVE, C && TG_COARSE_MEMBERS>(g, thr, cg, V, bc, out);
  }
}
#endif

// ------------------------------------------------------------------------------------------------
// ROLLED stage A / stage B for the edge kernel (-DTG_EDGE_ROLLED; a measured variant, NOT the default).  The unrolled
// forms above are ~60 KB of SASS; in the configuration kernel the 32 warps of a block walk that code together, but in the
// edge kernel every warp is at its own place in its own edge (ncu, profiles/r02a: stall_no_instruction 4.96 per issue,
// the largest stall).  Here the same probes are issued from loops over constant-memory tables: six spheres per
// iteration, one loop body for the seven arm links.  Parity-green, but slower: C2 0.663 -> 0.793 ms, C5 5.56 -> 6.16
// ms/step - the loop overhead and the serialised link bodies cost more than the instruction fetches save.
// ------------------------------------------------------------------------------------------------
__constant__ float TGM_ARM_OFF[7][3] = {
    {TGM_OFF_SHOULDER_PAN_X, TGM_OFF_SHOULDER_PAN_Y, TGM_OFF_SHOULDER_PAN_Z}, {TGM_OFF_SHOULDER_LIFT_X, TGM_OFF_SHOULDER_LIFT_Y, TGM_OFF_SHOULDER_LIFT_Z},
    {TGM_OFF_UPPERARM_ROLL_X, TGM_OFF_UPPERARM_ROLL_Y, TGM_OFF_UPPERARM_ROLL_Z}, {TGM_OFF_ELBOW_FLEX_X, TGM_OFF_ELBOW_FLEX_Y, TGM_OFF_ELBOW_FLEX_Z},
    {TGM_OFF_FOREARM_ROLL_X, TGM_OFF_FOREARM_ROLL_Y, TGM_OFF_FOREARM_ROLL_Z}, {TGM_OFF_WRIST_FLEX_X, TGM_OFF_WRIST_FLEX_Y, TGM_OFF_WRIST_FLEX_Z},
    {TGM_OFF_WRIST_ROLL_X, TGM_OFF_WRIST_ROLL_Y, TGM_OFF_WRIST_ROLL_Z}};
__constant__ int TGM_ARM_KIND[7] = {4, 3, 2, 3, 2, 3, 2};            // rotation axis of each arm joint: z, y, x, y, x, y, x
__constant__ int TGM_ARM_K0[8] = {55, 59, 66, 74, 80, 87, 93, 111};  // kernel-order sphere range of each arm frame

__device__ __forceinline__ void tg_classify_dyn(const TgThr& thr, int k, int code, TgClass& out) {
  out.collide |= code <= thr.coll_le[k];
  if (code < thr.free_ge[k]) {
    const uint32_t bit = 1u << (k & 31);
    const int w = k >> 5;
    out.amb[0] |= w == 0 ? bit : 0u;
    out.amb[1] |= w == 1 ? bit : 0u;
    out.amb[2] |= w == 2 ? bit : 0u;
    out.amb[3] |= w == 3 ? bit : 0u;
  }
}

// spheres [k0, k1) of a frame given by its voxel-space affine map, six per iteration
__device__ __forceinline__ void tg_group_rolled(const TgGrid& g, const TgThr& thr, const TgVox& V, int k0, int k1, TgClass& out) {
  for (int k = k0; k < k1 && !out.collide; k += 6) {
    int code[6];
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int kk = min(k + u, k1 - 1);
      const float4 l = TGM_K_SPH[kk];
      const float gx = fmaf(V.G[0], l.x, fmaf(V.G[1], l.y, fmaf(V.G[2], l.z, V.gt[0])));
      const float gy = fmaf(V.G[3], l.x, fmaf(V.G[4], l.y, fmaf(V.G[5], l.z, V.gt[1])));
      const float gz = fmaf(V.G[6], l.x, fmaf(V.G[7], l.y, fmaf(V.G[8], l.z, V.gt[2])));
      code[u] = tg_probe_if(g, k + u < k1, gx, gy, gz, TG_CODE_SKIPPED);
    }
#pragma unroll
    for (int u = 0; u < 6; ++u)
      if (k + u < k1) tg_classify_dyn(thr, k + u, code[u], out);
  }
}

// voxel-space map of a frame whose rotation is the base yaw only, translated by (tx, 0, tz) in base_link
__device__ __forceinline__ void tg_vox_yaw(const TgGrid& g, const TgPose& p, float tx, float tz, TgVox& V) {
  V.G[0] = p.cs; V.G[1] = -p.sn; V.G[2] = 0.f;
  V.G[3] = p.sn; V.G[4] = p.cs;  V.G[5] = 0.f;
  V.G[6] = 0.f;  V.G[7] = 0.f;   V.G[8] = g.vpm;
  V.gt[0] = fmaf(p.cs, tx, p.gx0);
  V.gt[1] = fmaf(p.sn, tx, p.gy0);
  V.gt[2] = g.vpm * tz;
}

__device__ __forceinline__ int tg_bound_rolled(const TgGrid& g, const TgVox& V, int b) {
  const float4 l = TGM_BOUND[b];
  return tg_probe(g, fmaf(V.G[0], l.x, fmaf(V.G[1], l.y, fmaf(V.G[2], l.z, V.gt[0]))),
                  fmaf(V.G[3], l.x, fmaf(V.G[4], l.y, fmaf(V.G[5], l.z, V.gt[1]))),
                  fmaf(V.G[6], l.x, fmaf(V.G[7], l.y, fmaf(V.G[8], l.z, V.gt[2]))));
}

__device__ __forceinline__ void tg_stage_a_rolled(const TgGrid& g, const TgThr& thr, const TgPose& p, TgClass& out) {
  TgVox V;
  tg_vox_yaw(g, p, 0.f, 0.f, V);
  tg_group_rolled(g, thr, V, 0, 3, out);  // the three big base spheres decide most collisions
  if (out.collide) return;
  tg_group_rolled(g, thr, V, 3, 20, out);
}

__device__ __forceinline__ void tg_stage_b_rolled(const TgGrid& g, const TgThr& thr, const TgPose& p, const float* q, TgClass& out) {
  const float tz = TGM_OFF_TORSO_Z + q[0];
  TgVox V;
  tg_vox_yaw(g, p, TGM_OFF_TORSO_X, tz, V);
  tg_group_rolled(g, thr, V, 20, 28, out);
  if (out.collide) return;
  {
    const int kb[4] = {28, 33, 44, 55};
#pragma unroll 1
    for (int b = 0; b < 3; ++b) {
      if (tg_bound_rolled(g, V, b) < thr.bound_ge[b]) {
        tg_group_rolled(g, thr, V, kb[b], kb[b + 1], out);
        if (out.collide) return;
      }
    }
  }
  TgFrame F, H;
  tg_frame_identity(F);
  F.t[0] = TGM_OFF_TORSO_X;
  F.t[2] = tz;
#pragma unroll 1
  for (int j = 0; j < 7; ++j) {
    const float ox = TGM_ARM_OFF[j][0], oy = TGM_ARM_OFF[j][1], oz = TGM_ARM_OFF[j][2];
    switch (TGM_ARM_KIND[j]) {  // the axis must be a compile-time constant inside tg_frame_child (register indexing)
      case 2: tg_frame_child(F, ox, oy, oz, 2, q[1 + j], H); break;
      case 3: tg_frame_child(F, ox, oy, oz, 3, q[1 + j], H); break;
      default: tg_frame_child(F, ox, oy, oz, 4, q[1 + j], H); break;
    }
    F = H;
    tg_vox_frame(g, p, F, V);
    if (tg_bound_rolled(g, V, 3 + j) < thr.bound_ge[3 + j]) {
      tg_group_rolled(g, thr, V, TGM_ARM_K0[j], TGM_ARM_K0[j + 1], out);
      if (out.collide) return;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// exact narrow phase, one lane per sphere (attachments, clearance output): walks the cell rows serially
// ------------------------------------------------------------------------------------------------
// true iff some cloud point p has |p - c|^2 <= R^2 (R already includes the point radius).
__device__ inline bool tg_sphere_hits_cloud(const TgGrid& g, float cx, float cy, float cz, float R) {
  const float R2 = R * R;
  const float Re = R + 1e-4f;  // cell-range margin: never skip a cell because of fp rounding
  const float inv = 1.0f / g.cpm;
  int z0 = max((int)floorf((cz - Re) * g.cpm) - g.cm[2], 0), z1 = min((int)floorf((cz + Re) * g.cpm) - g.cm[2], g.cnz - 1);
  for (int iz = z0; iz <= z1; ++iz) {
    const float zlo = (float)(iz + g.cm[2]) * inv;
    const float dz = fmaxf(0.f, fmaxf(zlo - cz, cz - (zlo + inv)));
    const float ry2 = Re * Re - dz * dz;
    if (ry2 < 0.f) continue;
    const float ry = sqrtf(ry2);
    int y0 = max((int)floorf((cy - ry) * g.cpm) - g.cm[1], 0), y1 = min((int)floorf((cy + ry) * g.cpm) - g.cm[1], g.cny - 1);
    for (int iy = y0; iy <= y1; ++iy) {
      const float ylo = (float)(iy + g.cm[1]) * inv;
      const float dy = fmaxf(0.f, fmaxf(ylo - cy, cy - (ylo + inv)));
      const float rx2 = ry2 - dy * dy;
      if (rx2 < 0.f) continue;
      const float rx = sqrtf(rx2) + 1e-4f;
      const int x0 = max((int)floorf((cx - rx) * g.cpmx) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpmx) - g.cm[0], g.cnx - 1);
      if (x0 > x1) continue;
      const int row = (iz * g.cny + iy) * g.cnx;
      const uint32_t s = __ldg(g.cell_start + row + x0), e = __ldg(g.cell_start + row + x1 + 1);
      for (uint32_t p = s; p < e; ++p) {
        const float4 P = __ldg(g.pts + p);
        const float dx = P.x - cx, dyy = P.y - cy, dzz = P.z - cz;
        if (fmaf(dx, dx, fmaf(dyy, dyy, dzz * dzz)) <= R2) return true;
      }
    }
  }
  return false;
}

// min over cloud points of (|p - c| - R), clipped to `limit` (search radius R + limit).
// `limit` may be negative (a deeper penetration is already known): R + limit must be positive.
__device__ inline float tg_sphere_clearance(const TgGrid& g, float cx, float cy, float cz, float R, float limit) {
  const float Rs = R + limit + 1e-4f;
  float best2 = (R + limit) * (R + limit);
  const float inv = 1.0f / g.cpm;
  int z0 = max((int)floorf((cz - Rs) * g.cpm) - g.cm[2], 0), z1 = min((int)floorf((cz + Rs) * g.cpm) - g.cm[2], g.cnz - 1);
  for (int iz = z0; iz <= z1; ++iz) {
    const float zlo = (float)(iz + g.cm[2]) * inv;
    const float dz = fmaxf(0.f, fmaxf(zlo - cz, cz - (zlo + inv)));
    const float ry2 = Rs * Rs - dz * dz;
    if (ry2 < 0.f || dz * dz > best2) continue;
    const float ry = sqrtf(ry2);
    int y0 = max((int)floorf((cy - ry) * g.cpm) - g.cm[1], 0), y1 = min((int)floorf((cy + ry) * g.cpm) - g.cm[1], g.cny - 1);
    for (int iy = y0; iy <= y1; ++iy) {
      const float ylo = (float)(iy + g.cm[1]) * inv;
      const float dy = fmaxf(0.f, fmaxf(ylo - cy, cy - (ylo + inv)));
      const float rx2 = ry2 - dy * dy;
      if (rx2 < 0.f || dz * dz + dy * dy > best2) continue;
      const float rx = sqrtf(rx2) + 1e-4f;
      const int x0 = max((int)floorf((cx - rx) * g.cpmx) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpmx) - g.cm[0], g.cnx - 1);
      if (x0 > x1) continue;
      const int row = (iz * g.cny + iy) * g.cnx;
      const uint32_t s = __ldg(g.cell_start + row + x0), e = __ldg(g.cell_start + row + x1 + 1);
      for (uint32_t p = s; p < e; ++p) {
        const float4 P = __ldg(g.pts + p);
        const float dx = P.x - cx, dyy = P.y - cy, dzz = P.z - cz;
        best2 = fminf(best2, fmaf(dx, dx, fmaf(dyy, dyy, dzz * dzz)));
      }
    }
  }
  return best2 < (R + limit) * (R + limit) ? sqrtf(best2) - R : limit;  // only a point that was really found changes the bound
}

// ------------------------------------------------------------------------------------------------
// exact narrow phase, one WARP per sphere (phase 2).  All 32 lanes must call it with the same
// arguments.  Lane r of a batch owns cell row (iz, iy) of the sphere's bounding box: it clips the
// row against the sphere (a chord in x) and reads the two cell_start entries that bracket the
// chord — the row's candidate points are one contiguous range because x is the fastest key.  The
// ranges of the 32 rows are then concatenated (warp prefix sum) and scanned together, 32 points per
// step with coalesced float4 loads; the first hit ends the walk through __any_sync.
// ------------------------------------------------------------------------------------------------

struct TgRows {  // cell rows of a sphere's bounding box
  int z0, y0, ny, nrows;
  float inv_ny;
};
__device__ __forceinline__ TgRows tg_rows(const TgGrid& g, float cy, float cz, float Re) {
  TgRows w;
  w.z0 = max((int)floorf((cz - Re) * g.cpm) - g.cm[2], 0);
  w.y0 = max((int)floorf((cy - Re) * g.cpm) - g.cm[1], 0);
  const int z1 = min((int)floorf((cz + Re) * g.cpm) - g.cm[2], g.cnz - 1);
  const int y1 = min((int)floorf((cy + Re) * g.cpm) - g.cm[1], g.cny - 1);
  w.ny = y1 - w.y0 + 1;
  const int nz = z1 - w.z0 + 1;
  w.nrows = (w.ny > 0 && nz > 0) ? w.ny * nz : 0;
  w.inv_ny = 1.0f / (float)max(w.ny, 1);
  return w;
}

// candidate point range [s, e) of row r (this lane's row): the chord of the sphere in that row
__device__ __forceinline__ void tg_row_range(const TgGrid& g, const TgRows& w, float cx, float cy, float cz, float Re, int r,
                                             uint32_t& s, uint32_t& e) {
  s = e = 0;
  if (r >= w.nrows) return;
  const float inv = 1.0f / g.cpm;
  const int jz = (int)(((float)r + 0.5f) * w.inv_ny);  // r / ny (exact: r < 2^20)
  const int iz = w.z0 + jz, iy = w.y0 + (r - jz * w.ny);
  const float zlo = (float)(iz + g.cm[2]) * inv, ylo = (float)(iy + g.cm[1]) * inv;
  const float dz = fmaxf(0.f, fmaxf(zlo - cz, cz - (zlo + inv)));
  const float dy = fmaxf(0.f, fmaxf(ylo - cy, cy - (ylo + inv)));
  const float rx2 = Re * Re - dz * dz - dy * dy;
  if (rx2 < 0.f) return;
  const float rx = sqrtf(rx2) + 1e-4f;
  const int x0 = max((int)floorf((cx - rx) * g.cpmx) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpmx) - g.cm[0], g.cnx - 1);
  if (x0 > x1) return;
  const int row = (iz * g.cny + iy) * g.cnx;
  s = __ldg(g.cell_start + row + x0);
  e = __ldg(g.cell_start + row + x1 + 1);
}

// scan the concatenation of the 32 lanes' ranges, 32 points per step; warp-uniform result
__device__ __forceinline__ bool tg_scan_ranges(const TgGrid& g, float cx, float cy, float cz, float R2, uint32_t s, uint32_t e) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t len = e - s;
  uint32_t incl = len;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  TG_STAT(4, (total + 31) / 32);
  TG_STAT(9, total);
  for (uint32_t j0 = 0; j0 < total; j0 += 32) {
    const uint32_t j = j0 + lane;
    int lo = 0;  // owner = first lane whose inclusive sum exceeds j (binary search over the warp's registers)
#pragma unroll
    for (int step = 16; step; step >>= 1) {
      const uint32_t v = __shfl_sync(0xffffffffu, incl, lo + step - 1);
      if (v <= j) lo += step;
    }
    lo = min(lo, 31);
    const uint32_t o_incl = __shfl_sync(0xffffffffu, incl, lo);
    const uint32_t o_e = __shfl_sync(0xffffffffu, e, lo);
    bool hit = false;
    if (j < total) {
      const float4 P = __ldg(g.pts + (o_e - (o_incl - j)));
      const float dx = P.x - cx, dyy = P.y - cy, dzz = P.z - cz;
      hit = fmaf(dx, dx, fmaf(dyy, dyy, dzz * dzz)) <= R2;
    }
    if (__any_sync(0xffffffffu, hit)) return true;
  }
  return false;
}

__device__ __forceinline__ bool tg_warp_sphere_hits(const TgGrid& g, float cx, float cy, float cz, float R) {
  const unsigned lane = threadIdx.x & 31;
  const float Re = R + 1e-4f;
  const TgRows w = tg_rows(g, cy, cz, Re);
  TG_STAT(2, 1);
  TG_STAT(R < 0.13f ? 5 : (R < 0.2f ? 6 : 7), 1);
  for (int r0 = 0; r0 < w.nrows; r0 += 32) {
    uint32_t s, e;
    TG_STAT(3, 1);
    tg_row_range(g, w, cx, cy, cz, Re, r0 + (int)lane, s, e);
    if (tg_scan_ranges(g, cx, cy, cz, R * R, s, e)) return true;
  }
  return false;
}

// Cooperative NEAREST point: min over cloud points within Rs of c of |p - c|^2, or Rs^2 when there is none, by a GROUP of G
// lanes (G = 4, 8, 16 or 32; 32 / G independent searches per warp, each group with its own arguments; `active` = false makes
// a group idle).  Same row ownership and coalesced scan as tg_warp_sphere_hits, with a running minimum instead of the early
// exit; z slices are visited centre-out (jc, jc + 1, jc - 1, ...): the nearest point is usually found in the first batch of
// rows, after which the radius - and with it the chords of the remaining rows, most of them then empty - shrinks.
// The searches are latency-bound (cell_start -> points, one dependent round trip per batch of rows), so what pays is
// searches in flight, not lanes per search.  Every lane of the warp must call it; loops run until the slowest group of the
// warp is done; the result is group-uniform.
template <int G>
__device__ __forceinline__ float tg_group_sphere_min_d2(const TgGrid& g, bool active, float cx, float cy, float cz, float Rs) {
  const unsigned full = 0xffffffffu;
  const int gl = (int)(threadIdx.x & (G - 1));
  float Re = Rs + 1e-4f;
  TgRows w = tg_rows(g, cy, cz, Re);
  if (!active) w.nrows = 0;
  float best = Rs * Rs;
  const int nz = w.ny > 0 ? w.nrows / w.ny : 0;
  const int jc = min(max((int)floorf(cz * g.cpm) - g.cm[2] - w.z0, 0), max(nz - 1, 0));
  const int mside = min(jc, nz - 1 - jc);
  const float inv = 1.0f / g.cpm;
  for (int r0 = 0; __any_sync(full, r0 < w.nrows); r0 += G) {
    uint32_t s = 0, e = 0;
    const int r = r0 + gl;
    if (r < w.nrows) {
      const int k = (int)(((float)r + 0.5f) * w.inv_ny);  // r / ny (exact: r < 2^20)
      const int jz = k <= 2 * mside ? jc + ((k & 1) ? (k + 1) / 2 : -(k / 2)) : (jc > nz - 1 - jc ? nz - 1 - k : k);
      const int iz = w.z0 + jz, iy = w.y0 + (r - k * w.ny);
      const float zlo = (float)(iz + g.cm[2]) * inv, ylo = (float)(iy + g.cm[1]) * inv;
      const float dz = fmaxf(0.f, fmaxf(zlo - cz, cz - (zlo + inv)));
      const float dy = fmaxf(0.f, fmaxf(ylo - cy, cy - (ylo + inv)));
      const float rx2 = Re * Re - dz * dz - dy * dy;
      if (rx2 >= 0.f) {
        const float rx = sqrtf(rx2) + 1e-4f;
        const int x0 = max((int)floorf((cx - rx) * g.cpmx) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpmx) - g.cm[0], g.cnx - 1);
        if (x0 <= x1) {
          const int row = (iz * g.cny + iy) * g.cnx;
          s = __ldg(g.cell_start + row + x0);
          e = __ldg(g.cell_start + row + x1 + 1);
        }
      }
    }
    uint32_t incl = e - s;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) {
      const uint32_t t = __shfl_up_sync(full, incl, o, G);
      if (gl >= o) incl += t;
    }
    const uint32_t total = __shfl_sync(full, incl, G - 1, G);
    if (!__any_sync(full, total != 0u)) continue;
    for (uint32_t j0 = 0; __any_sync(full, j0 < total); j0 += 2 * G) {  // two steps per iteration: both loads in flight together
      uint32_t idx[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const uint32_t j = j0 + (uint32_t)(u * G + gl);
        int lo = 0;  // owner row = first lane of the group whose inclusive sum exceeds j
#pragma unroll
        for (int step = G / 2; step; step >>= 1) {
          const uint32_t v = __shfl_sync(full, incl, lo + step - 1, G);
          if (v <= j) lo += step;
        }
        lo = min(lo, G - 1);
        const uint32_t o_incl = __shfl_sync(full, incl, lo, G);
        const uint32_t o_e = __shfl_sync(full, e, lo, G);
        idx[u] = j < total ? o_e - (o_incl - j) : 0xFFFFFFFFu;
      }
      float4 P[2];
#pragma unroll
      for (int u = 0; u < 2; ++u)
        if (idx[u] != 0xFFFFFFFFu) P[u] = __ldg(g.pts + idx[u]);
#pragma unroll
      for (int u = 0; u < 2; ++u)
        if (idx[u] != 0xFFFFFFFFu) {
          const float dx = P[u].x - cx, dyy = P[u].y - cy, dzz = P[u].z - cz;
          best = fminf(best, fmaf(dx, dx, fmaf(dyy, dyy, dzz * dzz)));
        }
    }
#pragma unroll
    for (int o = G / 2; o; o >>= 1) best = fminf(best, __shfl_xor_sync(full, best, o, G));
    Re = sqrtf(best) * 1.000001f + 1e-4f;  // never below the distance of the best point found so far
  }
#pragma unroll
  for (int o = G / 2; o; o >>= 1) best = fminf(best, __shfl_xor_sync(full, best, o, G));
  return best;
}

// Two spheres per warp: lanes 0-15 walk one sphere, lanes 16-31 another (each lane holds ITS half's sphere; `active`
// is half-uniform).  With the anisotropic cells a sphere touches <= 16 rows almost always, so a half warp is enough
// for the row lookups, and every instruction of the range arithmetic and of the scan now serves two spheres and two
// independent cell_start -> points load chains.  Same candidate sets, same fp32 test as tg_warp_sphere_hits.
// Returns the half-uniform verdict of this lane's sphere.
// Built with -DTG_HALFWARP only: parity-green but 6.5 % slower than one sphere per warp on the bench (the two halves
// wait for each other's scan steps, 16 points per step instead of 32); kept as a measured variant.
__device__ __forceinline__ bool tg_scan_ranges16(const TgGrid& g, float cx, float cy, float cz, float R2, uint32_t s, uint32_t e) {
  const unsigned lane = threadIdx.x & 31, hl = lane & 15u, sh = lane & 16u;
  uint32_t incl = e - s;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o, 16);
    if ((int)hl >= o) incl += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 15, 16);
  bool found = false;
  for (uint32_t j0 = 0;; j0 += 16) {
    const bool go = !found && j0 < total;  // half-uniform
    if (!__any_sync(0xffffffffu, go)) break;
    const uint32_t j = j0 + hl;
    int lo = 0;  // owner = first lane of the half whose inclusive sum exceeds j
#pragma unroll
    for (int step = 8; step; step >>= 1) {
      const uint32_t v = __shfl_sync(0xffffffffu, incl, lo + step - 1, 16);
      if (v <= j) lo += step;
    }
    lo = min(lo, 15);
    const uint32_t o_incl = __shfl_sync(0xffffffffu, incl, lo, 16);
    const uint32_t o_e = __shfl_sync(0xffffffffu, e, lo, 16);
    bool hit = false;
    if (go && j < total) {
      const float4 P = __ldg(g.pts + (o_e - (o_incl - j)));
      const float dx = P.x - cx, dyy = P.y - cy, dzz = P.z - cz;
      hit = fmaf(dx, dx, fmaf(dyy, dyy, dzz * dzz)) <= R2;
    }
    found |= ((__ballot_sync(0xffffffffu, hit) >> sh) & 0xFFFFu) != 0u;
  }
  return found;
}

__device__ __forceinline__ bool tg_halfwarp_sphere_hits(const TgGrid& g, bool active, float cx, float cy, float cz, float R) {
  const unsigned hl = threadIdx.x & 15u;
  const float Re = R + 1e-4f;
  TgRows w = tg_rows(g, cy, cz, Re);
  if (!active) w.nrows = 0;
  if (active && hl == 0) {
    TG_STAT_LANE(2, 1);
    TG_STAT_LANE(R < 0.13f ? 5 : (R < 0.2f ? 6 : 7), 1);
  }
  bool found = false;
  for (int r0 = 0;; r0 += 16) {
    const bool more = !found && r0 < w.nrows;  // half-uniform
    if (!__any_sync(0xffffffffu, more)) break;
    uint32_t s = 0, e = 0;
    if (more) tg_row_range(g, w, cx, cy, cz, Re, r0 + (int)hl, s, e);
    found |= tg_scan_ranges16(g, cx, cy, cz, R * R, s, e);
  }
  return found;
}

// World centre (metres) + radius of kernel-order sphere k for the planner model (head = 0).
__device__ __forceinline__ float4 tg_sphere_world(const TgFrame* L, const TgBase& b, int k) {
  const int j = TGM_K_REF[k];
  const float4 l = TGM_SPH_REF[j];
  const float3 pb = tg_frame_apply(L[TGM_SPH_LINK[j]], l.x, l.y, l.z);
  const float3 pw = tg_to_world(b, pb.x, pb.y, pb.z);
  return make_float4(pw.x, pw.y, pw.z, l.w);
}

// Optional self-collision term (flag TG_FLAG_SELF, off by default): sphere-vs-sphere between the link pairs of the
// reference's SELF_COLLISION_PAIRS (TyGrit/rl/train.py:78-107; VAMP's own pair list is unavailable).  Evaluated in the
// base_link frame (the base pose cancels), head = 0.  Link pairs whose enclosing spheres are apart are pruned, then
// sphere i of link a is tested against link b's enclosing sphere before its spheres.  Returns
// min(|c_i - c_j| - r_i - r_j) over the tested pairs, clipped from above at `stop_above` pruning (pairs that cannot
// come below it are skipped); with stop_at_hit the walk ends at the first touching pair (returns <= 0).
static __device__ __noinline__ float tg_self_clearance(const float* q8, float stop_above, bool stop_at_hit) {
  float q10[10];
#pragma unroll
  for (int i = 0; i < 8; ++i) q10[i] = q8[i];
  q10[8] = q10[9] = 0.f;
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q10, L);
  float best = stop_above;
  for (int a = 0; a < TGM_N_LINKS; ++a) {
    const unsigned partners = TGM_SELF_PARTNERS[a];
    const int na = TGM_LINK_SPH_COUNT[a];
    if (partners == 0u || na == 0) continue;
    const float4 ba = TGM_LINK_BOUND[a];
    const float3 ca = tg_frame_apply(L[a], ba.x, ba.y, ba.z);
    for (int b = 0; b < TGM_N_LINKS; ++b) {
      const int nb = TGM_LINK_SPH_COUNT[b];
      if (!((partners >> b) & 1u) || nb == 0) continue;
      const float4 bb = TGM_LINK_BOUND[b];
      const float3 cb = tg_frame_apply(L[b], bb.x, bb.y, bb.z);
      {
        const float dx = ca.x - cb.x, dy = ca.y - cb.y, dz = ca.z - cb.z;
        const float reach = ba.w + bb.w + fmaxf(best, 0.f);
        if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) > reach * reach) continue;
      }
      for (int i = 0; i < na; ++i) {
        const float4 si = TGM_SPH_REF[TGM_LINK_SPH_FIRST[a] + i];
        const float3 ci = tg_frame_apply(L[a], si.x, si.y, si.z);
        {
          const float dx = ci.x - cb.x, dy = ci.y - cb.y, dz = ci.z - cb.z;
          const float reach = si.w + bb.w + fmaxf(best, 0.f);
          if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) > reach * reach) continue;
        }
        for (int j = 0; j < nb; ++j) {
          const float4 sj = TGM_SPH_REF[TGM_LINK_SPH_FIRST[b] + j];
          const float3 cj = tg_frame_apply(L[b], sj.x, sj.y, sj.z);
          const float dx = ci.x - cj.x, dy = ci.y - cj.y, dz = ci.z - cj.z;
          const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz)), rr = si.w + sj.w;
          if (stop_at_hit) {
            if (d2 <= rr * rr) return -1.f;
          } else {
            best = fminf(best, sqrtf(d2) - rr);
          }
        }
      }
    }
  }
  return best;
}
__device__ __forceinline__ bool tg_self_hit(const float* q8) { return tg_self_clearance(q8, 0.f, true) < 0.f; }

// Sphere obstacles and end-effector attachment (vamp_preview.py:242-304), brute force, exact.
static __device__ __noinline__ bool tg_extras_hit(const TgGrid& g, const TgExtras& ex, const float* cfg /*x,y,th,q0..q7*/) {
  float q10[10];
#pragma unroll
  for (int i = 0; i < 8; ++i) q10[i] = cfg[3 + i];
  q10[8] = q10[9] = 0.f;
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q10, L);
  const TgBase b = tg_base(cfg[0], cfg[1], cfg[2]);
  for (int k = 0; k < TG_N_SPHERES && ex.n_obs > 0; ++k) {
    const float4 c = tg_sphere_world(L, b, k);
    for (int o = 0; o < ex.n_obs; ++o) {
      const float4 ob = ex.obs[o];
      const float dx = c.x - ob.x, dy = c.y - ob.y, dz = c.z - ob.z, R = c.w + ob.w;
      if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= R * R) return true;
    }
  }
  for (int a = 0; a < ex.n_att; ++a) {
    const float4 l = ex.att[a];
    const float3 pb = tg_frame_apply(L[TGM_GRIPPER_LINK], l.x, l.y, l.z);
    const float3 c = tg_to_world(b, pb.x, pb.y, pb.z);
    for (int o = 0; o < ex.n_obs; ++o) {
      const float4 ob = ex.obs[o];
      const float dx = c.x - ob.x, dy = c.y - ob.y, dz = c.z - ob.z, R = l.w + ob.w;
      if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= R * R) return true;
    }
    if (!g.empty && tg_sphere_hits_cloud(g, c.x, c.y, c.z, l.w + g.r_pt)) return true;
  }
  return false;
}

// Resolve an item with one lane (queue-overflow fallback): exact test of every ambiguous sphere.
static __device__ __noinline__ bool tg_resolve_item(const TgGrid& g, const float* cfg, const uint32_t* mask) {
  float q10[10];
#pragma unroll
  for (int i = 0; i < 8; ++i) q10[i] = cfg[3 + i];
  q10[8] = q10[9] = 0.f;
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q10, L);
  const TgBase b = tg_base(cfg[0], cfg[1], cfg[2]);
  for (int w = 0; w < 4; ++w) {
    uint32_t m = mask[w];
    while (m) {
      const int bit = __ffs(m) - 1;
      m &= m - 1;
      const float4 c = tg_sphere_world(L, b, 32 * w + bit);
      if (tg_sphere_hits_cloud(g, c.x, c.y, c.z, c.w + g.r_pt)) return true;
    }
  }
  return false;
}

// Call-by-copy fronts of the two out-of-line helpers.  A pointer argument to a __noinline__ function forces the caller's
// array into local memory for its whole life: with the configuration (x, y, theta, q0..q7) living on the stack, every use of
// a joint angle in stage B was a local-memory load behind the store that put it there.  Passing a copy at these (rare) call
// sites keeps the hot copy in registers: phase 1 - 6 % (3.524 -> 3.311 ms per 64 Mi configurations, free-heavy batch
// 8.43 -> 7.92 ms; profiles/r03q_cfg_copy.txt).
__device__ __forceinline__ bool tg_extras_hit_v(const TgGrid& g, const TgExtras& ex, const float* cfg) {
  float c2[11];
#pragma unroll
  for (int i = 0; i < 11; ++i) c2[i] = cfg[i];
  return tg_extras_hit(g, ex, c2);
}
__device__ __forceinline__ bool tg_resolve_item_v(const TgGrid& g, const float* cfg, const uint32_t* mask) {
  float c2[11];
  uint32_t m2[4];
#pragma unroll
  for (int i = 0; i < 11; ++i) c2[i] = cfg[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) m2[i] = mask[i];
  return tg_resolve_item(g, c2, m2);
}
