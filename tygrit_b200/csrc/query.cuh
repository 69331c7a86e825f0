// query.cuh — device functions of the fused FK + sphere-query path (K2 + K3).
//
// Phase 1 (tg_classify): one lane = one configuration.  The planner's robot model has the head
// fixed, so 55 of the 111 spheres (base, torso, head) only need the SE(2) base pose and the torso
// lift; the remaining 56 ride the 7-joint arm chain.  Every sphere centre is produced directly in
// distance-grid coordinates (voxels) and costs ONE byte load from the L2-resident grid, giving a
// three-way answer per sphere: definitely free / definitely colliding / ambiguous.
// Phase 2 (tg_sphere_hits_cloud): exact test of an ambiguous sphere against the sorted cell rows.
#pragma once
#include "common.cuh"
#include "fk.cuh"

__device__ __forceinline__ int tg_probe(const TgGrid& g, float gx, float gy, float gz) {
  const float fx = fminf(fmaxf(gx + g.bias[0], TG_MAGIC), g.hi[0]);
  const float fy = fminf(fmaxf(gy + g.bias[1], TG_MAGIC), g.hi[1]);
  const float fz = fminf(fmaxf(gz + g.bias[2], TG_MAGIC), g.hi[2]);
  const unsigned idx = __float_as_uint(fx) + (unsigned)g.nx * __float_as_uint(fy) + (unsigned)g.nxy * __float_as_uint(fz) - g.idx_bias;
  return (int)__ldg(g.df + idx);
}

struct TgClass {
  bool collide;
  uint32_t amb[4];
};

#define TG_CLASSIFY_ONE(k, code)                                 \
  do {                                                           \
    const int _c = (code);                                       \
    if (_c <= thr.coll_le[k]) out.collide = true;                \
    else if (_c < thr.free_ge[k]) out.amb[(k) >> 5] |= 1u << ((k) & 31); \
  } while (0)

// spheres [K0, K1) of a frame whose rotation is the base yaw only:  p_base = l + (tx, 0, tz)
template <int K0, int K1>
__device__ __forceinline__ void tg_probe_yaw_frame(const TgGrid& g, const TgThr& thr, float cs, float sn, float gtx, float gty,
                                                   float gtz, TgClass& out) {
#pragma unroll 5
  for (int k = K0; k < K1; ++k) {
    const float4 l = TGM_K_SPH[k];
    const float gx = fmaf(cs, l.x, fmaf(-sn, l.y, gtx));
    const float gy = fmaf(sn, l.x, fmaf(cs, l.y, gty));
    const float gz = fmaf(g.vpm, l.z, gtz);
    TG_CLASSIFY_ONE(k, tg_probe(g, gx, gy, gz));
  }
}

// spheres [K0, K1) of an arm frame F (base_link frame) under base pose (cs, sn, gtx0, gty0) in voxel units
template <int K0, int K1>
__device__ __forceinline__ void tg_probe_arm_frame(const TgGrid& g, const TgThr& thr, const TgFrame& F, float cs, float sn,
                                                   float gtx0, float gty0, TgClass& out) {
  // G = vpm * Rz(theta) * R ;  gt = vpm * (Rz(theta) t + p)
  float G[9], gt[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    G[c] = fmaf(cs, F.r[c], -sn * F.r[3 + c]);
    G[3 + c] = fmaf(sn, F.r[c], cs * F.r[3 + c]);
    G[6 + c] = g.vpm * F.r[6 + c];
  }
  gt[0] = fmaf(cs, F.t[0], fmaf(-sn, F.t[1], gtx0));
  gt[1] = fmaf(sn, F.t[0], fmaf(cs, F.t[1], gty0));
  gt[2] = g.vpm * F.t[2];
#pragma unroll 4
  for (int k = K0; k < K1; ++k) {
    const float4 l = TGM_K_SPH[k];
    const float gx = fmaf(G[0], l.x, fmaf(G[1], l.y, fmaf(G[2], l.z, gt[0])));
    const float gy = fmaf(G[3], l.x, fmaf(G[4], l.y, fmaf(G[5], l.z, gt[1])));
    const float gz = fmaf(G[6], l.x, fmaf(G[7], l.y, fmaf(G[8], l.z, gt[2])));
    TG_CLASSIFY_ONE(k, tg_probe(g, gx, gy, gz));
  }
}

// Three-way classification of one configuration against the distance-bound grid.
// q: 8 arm joints (torso first); base pose (bx, by, bth).
__device__ __forceinline__ void tg_classify(const TgGrid& g, const TgThr& thr, const float* q, float bx, float by, float bth,
                                            TgClass& out) {
  out.collide = false;
  out.amb[0] = out.amb[1] = out.amb[2] = out.amb[3] = 0u;
  float s, c;
  sincosf(bth, &s, &c);
  const float cs = g.vpm * c, sn = g.vpm * s;
  const float gtx0 = g.vpm * bx, gty0 = g.vpm * by;
  // frame 0: base_link + torso_fixed
  tg_probe_yaw_frame<0, 20>(g, thr, cs, sn, gtx0, gty0, 0.f, out);
  // frame 1: torso_lift + head (head joints fixed at 0)
  const float tz = TGM_OFF_TORSO_Z + q[0];
  {
    const float gtx1 = fmaf(cs, TGM_OFF_TORSO_X, gtx0), gty1 = fmaf(sn, TGM_OFF_TORSO_X, gty0);
    tg_probe_yaw_frame<20, 55>(g, thr, cs, sn, gtx1, gty1, g.vpm * tz, out);
  }
  if (out.collide) return;
  // arm chain in base_link
  TgFrame T, F, H;
  tg_frame_identity(T);
  T.t[0] = TGM_OFF_TORSO_X;
  T.t[2] = tz;
  tg_frame_child(T, TGM_OFF_SHOULDER_PAN_X, TGM_OFF_SHOULDER_PAN_Y, TGM_OFF_SHOULDER_PAN_Z, 4, q[1], F);
  tg_probe_arm_frame<55, 59>(g, thr, F, cs, sn, gtx0, gty0, out);
  tg_frame_child(F, TGM_OFF_SHOULDER_LIFT_X, TGM_OFF_SHOULDER_LIFT_Y, TGM_OFF_SHOULDER_LIFT_Z, 3, q[2], H);
  tg_probe_arm_frame<59, 66>(g, thr, H, cs, sn, gtx0, gty0, out);
  tg_frame_child(H, TGM_OFF_UPPERARM_ROLL_X, TGM_OFF_UPPERARM_ROLL_Y, TGM_OFF_UPPERARM_ROLL_Z, 2, q[3], F);
  tg_probe_arm_frame<66, 74>(g, thr, F, cs, sn, gtx0, gty0, out);
  if (out.collide) return;
  tg_frame_child(F, TGM_OFF_ELBOW_FLEX_X, TGM_OFF_ELBOW_FLEX_Y, TGM_OFF_ELBOW_FLEX_Z, 3, q[4], H);
  tg_probe_arm_frame<74, 80>(g, thr, H, cs, sn, gtx0, gty0, out);
  tg_frame_child(H, TGM_OFF_FOREARM_ROLL_X, TGM_OFF_FOREARM_ROLL_Y, TGM_OFF_FOREARM_ROLL_Z, 2, q[5], F);
  tg_probe_arm_frame<80, 87>(g, thr, F, cs, sn, gtx0, gty0, out);
  if (out.collide) return;
  tg_frame_child(F, TGM_OFF_WRIST_FLEX_X, TGM_OFF_WRIST_FLEX_Y, TGM_OFF_WRIST_FLEX_Z, 3, q[6], H);
  tg_probe_arm_frame<87, 93>(g, thr, H, cs, sn, gtx0, gty0, out);
  tg_frame_child(H, TGM_OFF_WRIST_ROLL_X, TGM_OFF_WRIST_ROLL_Y, TGM_OFF_WRIST_ROLL_Z, 2, q[7], F);
  tg_probe_arm_frame<93, 111>(g, thr, F, cs, sn, gtx0, gty0, out);
}

// ------------------------------------------------------------------------------------------------
// exact narrow phase
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
      const int x0 = max((int)floorf((cx - rx) * g.cpm) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpm) - g.cm[0], g.cnx - 1);
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
      const int x0 = max((int)floorf((cx - rx) * g.cpm) - g.cm[0], 0), x1 = min((int)floorf((cx + rx) * g.cpm) - g.cm[0], g.cnx - 1);
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
  return sqrtf(best2) - R;
}

// World centre (metres) + radius of kernel-order sphere k for the planner model (head = 0).
__device__ __forceinline__ float4 tg_sphere_world(const TgFrame* L, const TgBase& b, int k) {
  const int j = TGM_K_REF[k];
  const float4 l = TGM_SPH_REF[j];
  const float3 pb = tg_frame_apply(L[TGM_SPH_LINK[j]], l.x, l.y, l.z);
  const float3 pw = tg_to_world(b, pb.x, pb.y, pb.z);
  return make_float4(pw.x, pw.y, pw.z, l.w);
}

// Sphere obstacles and end-effector attachment (vamp_preview.py:242-304), brute force, exact.
__device__ inline bool tg_extras_hit(const TgGrid& g, const TgExtras& ex, const float* cfg /*x,y,th,q0..q7*/) {
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

// Resolve an item: exact test of every ambiguous sphere; true = collision.
__device__ inline bool tg_resolve_item(const TgGrid& g, const float* cfg, const uint32_t* mask, unsigned long long* n_pts) {
  float q10[10];
#pragma unroll
  for (int i = 0; i < 8; ++i) q10[i] = cfg[3 + i];
  q10[8] = q10[9] = 0.f;
  TgFrame L[TGM_N_LINKS];
  tg_fk_links(q10, L);
  const TgBase b = tg_base(cfg[0], cfg[1], cfg[2]);
  (void)n_pts;
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
