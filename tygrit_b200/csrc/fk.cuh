// fk.cuh — Fetch forward kinematics device functions (fp32, FMA, accurate sincosf).
//
// Chain, joint axes and offsets follow TyGrit/kinematics/fetch/fk_numpy.py:65-155; the base pose
// composition follows TyGrit/utils/transforms.py:116-127.  The chain is evaluated in the
// base_link frame (|t| < 1.5 m, so fp32 rounding stays ~1e-7 m) and moved to the world by the
// SE(2) base pose in a single last step; this keeps sphere centres within 1e-6 m of the fp64
// oracle for workspaces of a few metres.
#pragma once
#include "common.cuh"
#include "fetch_model.cuh"

// sin/cos with a short inline fast path (|x| < 1000: three-term Cody-Waite reduction by pi/2 and
// degree-7/8 minimax polynomials, <= 1.5 ulp, checked against fp64 in tests/test_oracle.py) and ONE
// shared out-of-line slow path.  The library sincosf inlines its Payne-Hanek branch at every call
// site; with eight joints per configuration that was a quarter of the query kernels' code and
// showed up as instruction-cache stalls.
// (The slow path returns by value: handing the caller's result pointers to an out-of-line function would force every sine and
// cosine of the kernels through local memory.)
static __device__ __noinline__ float2 tg_sincos_slow(float x) {
  float s, c;
  sincosf(x, &s, &c);
  return make_float2(s, c);
}

__device__ __forceinline__ void tg_sincos(float x, float* sn, float* cs) {
  if (!(fabsf(x) < 1000.0f)) {
    const float2 r = tg_sincos_slow(x);
    *sn = r.x;
    *cs = r.y;
    return;
  }
  const float k = rintf(x * 0.636619747f);
  float r = fmaf(k, -1.57079601e+00f, x);
  r = fmaf(k, -3.13916473e-07f, r);
  r = fmaf(k, -5.39030253e-15f, r);
  const float s = r * r;
  float z = 2.86567956e-6f;
  z = fmaf(z, s, -1.98559923e-4f);
  z = fmaf(z, s, 8.33338592e-3f);
  z = fmaf(z, s, -1.66666672e-1f);
  const float ps = fmaf(z, r * s, r);
  z = 2.44677067e-5f;
  z = fmaf(z, s, -1.38877297e-3f);
  z = fmaf(z, s, 4.16666567e-2f);
  z = fmaf(z, s, -5.0e-1f);
  const float pc = fmaf(z, s, 1.0f);
  const int i = (int)k;
  float a = (i & 1) ? pc : ps, b = (i & 1) ? ps : pc;
  if (i & 2) a = -a;
  if ((i + 1) & 2) b = -b;
  *sn = a;
  *cs = b;
}

struct TgFrame {
  float r[9];  // row-major rotation
  float t[3];
};

__device__ __forceinline__ void tg_frame_identity(TgFrame& F) {
  F.r[0] = 1.f; F.r[1] = 0.f; F.r[2] = 0.f;
  F.r[3] = 0.f; F.r[4] = 1.f; F.r[5] = 0.f;
  F.r[6] = 0.f; F.r[7] = 0.f; F.r[8] = 1.f;
  F.t[0] = F.t[1] = F.t[2] = 0.f;
}

// C = P * Trans(o) * Rot_axis(angle);  kind: 0 fixed, 1 prismatic z (o.z += q), 2/3/4 = x/y/z
__device__ __forceinline__ void tg_frame_child(const TgFrame& P, float ox, float oy, float oz, int kind,
                                               float q, TgFrame& C) {
  if (kind == 1) oz += q;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    C.t[i] = fmaf(P.r[3 * i + 0], ox, fmaf(P.r[3 * i + 1], oy, fmaf(P.r[3 * i + 2], oz, P.t[i])));
  if (kind < 2) {
#pragma unroll
    for (int i = 0; i < 9; ++i) C.r[i] = P.r[i];
    return;
  }
  float s, c;
  tg_sincos(q, &s, &c);
  // columns (a, b) that mix: x-axis -> (1,2), y-axis -> (2,0), z-axis -> (0,1)
  const int a = (kind == 2) ? 1 : (kind == 3) ? 2 : 0;
  const int b = (kind == 2) ? 2 : (kind == 3) ? 0 : 1;
  const int k = 3 - a - b;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float pa = P.r[3 * i + a], pb = P.r[3 * i + b];
    C.r[3 * i + a] = fmaf(pa, c, pb * s);     // new col a =  c*a + s*b
    C.r[3 * i + b] = fmaf(pb, c, -(pa * s));  // new col b = -s*a + c*b
    C.r[3 * i + k] = P.r[3 * i + k];
  }
}

// All 15 link frames in base_link for q[10] (order = fetch_model LINKS).
__device__ __forceinline__ void tg_fk_links(const float* q, TgFrame* F) {
  tg_frame_identity(F[0]);
#pragma unroll
  for (int i = 1; i < TGM_N_LINKS; ++i) {
    const int kind = TGM_LINK_KIND[i];
    const int slot = TGM_LINK_QSLOT[i];
    const float qi = slot >= 0 ? q[slot] : 0.f;
    tg_frame_child(F[TGM_LINK_PARENT[i]], TGM_LINK_OFF[i][0], TGM_LINK_OFF[i][1], TGM_LINK_OFF[i][2], kind,
                   qi, F[i]);
  }
}

struct TgBase {  // SE(2) base pose
  float c, s, x, y;
};
__device__ __forceinline__ TgBase tg_base(float x, float y, float theta) {
  TgBase b;
  tg_sincos(theta, &b.s, &b.c);
  b.x = x;
  b.y = y;
  return b;
}

// world position of a base-frame point
__device__ __forceinline__ float3 tg_to_world(const TgBase& b, float px, float py, float pz) {
  return make_float3(fmaf(b.c, px, fmaf(-b.s, py, b.x)), fmaf(b.s, px, fmaf(b.c, py, b.y)), pz);
}

// base-frame centre of a sphere given in a link frame
__device__ __forceinline__ float3 tg_frame_apply(const TgFrame& F, float lx, float ly, float lz) {
  return make_float3(fmaf(F.r[0], lx, fmaf(F.r[1], ly, fmaf(F.r[2], lz, F.t[0]))),
                     fmaf(F.r[3], lx, fmaf(F.r[4], ly, fmaf(F.r[5], lz, F.t[1]))),
                     fmaf(F.r[6], lx, fmaf(F.r[7], ly, fmaf(F.r[8], lz, F.t[2]))));
}
