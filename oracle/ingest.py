"""Robot self-filter and depth -> world-cloud ingest, restated on the CPU.

Oracle (test infrastructure).  Follows
* ``TyGrit/kinematics/robot_filter.py:13-64``  (self-filter),
* ``TyGrit/utils/depth.py:13-84``              (depth back-projection),
* ``TyGrit/utils/pointcloud.py:13-90``         (voxel dedup, merge, crop, ground).

Numeric conventions that matter for bit-exactness (all checked against the
reference's own functions by ``tools/make_golden.py`` under numpy 2.x, the
numpy of this image; the reference pins numpy<2 where ``f32_array - f64_scalar``
stays f32 — the difference is below 1 ulp of the result and inside the exempt
band):
* self-filter: centres are computed in fp64 and cast to f32
  (``robot_filter.py:43-49``); ``d2 = ((dx*dx + dy*dy) + dz*dz)`` in f32 with no
  fused multiply-add; a point is REMOVED iff ``d2 < f32(r*r)`` (strict).
* depth: pixel index (f32) minus ``K[.,2]`` (np.float64 scalar) promotes to f64,
  so ``x = ((u - cx) * d) / fx`` is evaluated in f64 and rounded to f32 once
  (``depth.py:41-49``); the world transform ``p @ R.T + t`` is f32
  (``depth.py:79-83``).
"""

from __future__ import annotations

import numpy as np

from . import fetch_model as fm
from . import fk


def filter_centres_f32(q10, base_xyt) -> np.ndarray:
    """(B,111,3) f32 sphere centres for the self-filter (10-DOF FK incl. head)."""
    q = np.atleast_2d(np.asarray(q10, dtype=np.float64))
    c = fk.sphere_centres(q[:, :8], base_xyt, q[:, 8:10])
    return c.astype(np.float32)


def self_filter_keep(points_f32: np.ndarray, centres_f32: np.ndarray, radius: float = 0.08) -> np.ndarray:
    """keep[i] = not any_s( d2(i, s) < r^2 ), f32 arithmetic as the reference."""
    p = np.asarray(points_f32, dtype=np.float32).reshape(-1, 3)
    c = np.asarray(centres_f32, dtype=np.float32).reshape(-1, 3)
    r2 = np.float32(radius * radius)
    keep = np.ones(p.shape[0], dtype=bool)
    for s in range(c.shape[0]):
        dx = p[:, 0] - c[s, 0]
        dy = p[:, 1] - c[s, 1]
        dz = p[:, 2] - c[s, 2]
        d2 = (dx * dx + dy * dy) + dz * dz
        keep &= ~(d2 < r2)
    return keep


def self_filter_signed_band(points, centres_f64, radius: float = 0.08) -> np.ndarray:
    """fp64 ``min_s |p - c_s| - r`` per point: the exemption band is |value| <= 1e-5."""
    p = np.asarray(points, dtype=np.float64).reshape(-1, 3)
    c = np.asarray(centres_f64, dtype=np.float64).reshape(-1, 3)
    best = np.full(p.shape[0], np.inf)
    for s in range(c.shape[0]):
        best = np.minimum(best, np.linalg.norm(p - c[s], axis=1))
    return best - radius


def depth_to_camera_points(depth: np.ndarray, K: np.ndarray, stride: int = 1):
    """Camera-frame points of the stride-decimated image, plus the pixel validity mask.

    Returns ``(xyz (H',W',3) f32, valid (H',W') bool)`` where ``valid = depth > 0``
    (``depth.py:32-49``).  Row-major order of ``xyz[valid]`` equals the
    reference's ``np.where`` order.
    """
    d = np.nan_to_num(np.asarray(depth, dtype=np.float32))
    K = np.asarray(K, dtype=np.float64)
    if stride > 1:
        d = d[::stride, ::stride]
        K = K.copy()
        K[0, :] /= stride
        K[1, :] /= stride
    h, w = d.shape
    u = np.arange(w, dtype=np.float32)[None, :].astype(np.float64)
    v = np.arange(h, dtype=np.float32)[:, None].astype(np.float64)
    dd = d.astype(np.float64)
    x = ((u - K[0, 2]) * dd / K[0, 0]).astype(np.float32)
    y = ((v - K[1, 2]) * dd / K[1, 1]).astype(np.float32)
    xyz = np.stack([x, y, d], axis=-1)
    return xyz, d > 0


def depth_to_world(depth, K, T_wc, z_range=(0.2, 3.0), stride: int = 4):
    """World-frame cloud exactly as ``depth_to_world_pointcloud`` (depth.py:52-84).

    Returns ``(points (N,3) f32, keep (H',W') bool)``.
    """
    xyz, valid = depth_to_camera_points(depth, K, stride)
    z = xyz[..., 2]
    keep = valid & (z >= np.float32(z_range[0])) & (z <= np.float32(z_range[1]))
    pc = xyz[keep]
    T = np.asarray(T_wc, dtype=np.float64)
    rot = T[:3, :3].astype(np.float32)
    t = T[:3, 3].astype(np.float32)
    # explicit f32 sum in x,y,z order (BLAS sgemm order is unspecified; the
    # difference is <= 1 ulp and covered by the tolerance in the tests)
    pw = (pc[:, 0:1] * rot[None, :, 0] + pc[:, 1:2] * rot[None, :, 1]) + pc[:, 2:3] * rot[None, :, 2] + t
    return pw.astype(np.float32), keep


def crop_xy(points, center, radius):
    """XY-cylinder crop, inclusive (pointcloud.py:67-80)."""
    p = np.asarray(points, dtype=np.float32)
    c = np.asarray(center, dtype=np.float32)
    d = p[:, :2] - c[:2]
    d2 = np.sum(d * d, axis=1)
    return d2 <= radius * radius


def ground_keep(points, z_thr):
    """z > thr (pointcloud.py:83-90)."""
    return np.asarray(points)[:, 2] > z_thr


def voxel_keys(points, q):
    """floor(p / q) int64 with the division in f32 (pointcloud.py:26)."""
    return np.floor(np.asarray(points, dtype=np.float32) / np.float32(q)).astype(np.int64)


def voxel_first_index(points, voxel_size):
    """Indices (ascending) of the first point of each voxel (pointcloud.py:13-31)."""
    keys = voxel_keys(points, max(voxel_size, 1e-9))
    _, idx = np.unique(keys, axis=0, return_index=True)
    idx.sort()
    return idx


def merge_add_index(base, add, radius):
    """Indices into ``add`` kept by ``merge_dedup`` (pointcloud.py:34-64)."""
    q = max(radius, 1e-6)
    allk = np.vstack((voxel_keys(base, q), voxel_keys(add, q)))
    _, ui = np.unique(allk, axis=0, return_index=True)
    ai = ui[ui >= len(base)] - len(base)
    ai.sort()
    return ai


def _to_camera_f32(points, T_wc):
    """``(points - t) @ rot`` in f32, explicit x, y, z summation order (BLAS order is unspecified; <= 1 ulp)."""
    T = np.asarray(T_wc, dtype=np.float64)
    rot = T[:3, :3].astype(np.float32)
    t = T[:3, 3].astype(np.float32)
    d = np.asarray(points, dtype=np.float32) - t
    return (d[:, 0:1] * rot[None, 0, :] + d[:, 1:2] * rot[None, 1, :]) + d[:, 2:3] * rot[None, 2, :]


def frustum_mask(points, K, T_wc, z_range, width=640, height=480):
    """``points_in_frustum_mask`` (utils/pointcloud.py:93-133): True = inside the camera frustum."""
    K = np.asarray(K, dtype=np.float64)
    pc = _to_camera_f32(points, T_wc)
    z = pc[:, 2]
    mz = (z >= np.float32(z_range[0])) & (z <= np.float32(z_range[1]))
    zs = np.where(mz, z, np.float32(1.0)).astype(np.float64)
    u = pc[:, 0].astype(np.float64) * K[0, 0] / zs + K[0, 2]
    v = pc[:, 1].astype(np.float64) * K[1, 1] / zs + K[1, 2]
    return mz & (u >= 0) & (u < width) & (v >= 0) & (v < height)


def raycast_keep(points, depth, K, T_wc, merge_radius):
    """Free-space carve of ``_update_ray_casting`` (pointcloud_scene.py:218-262): False = removed."""
    K = np.asarray(K, dtype=np.float64)
    depth = np.asarray(depth, dtype=np.float32)
    pc = _to_camera_f32(points, T_wc)
    zc = pc[:, 2]
    valid = zc > 1e-3
    zs = np.where(valid, zc, np.float32(1.0)).astype(np.float64)
    u = pc[:, 0].astype(np.float64) * K[0, 0] / zs + K[0, 2]
    v = pc[:, 1].astype(np.float64) * K[1, 1] / zs + K[1, 2]
    ui, vi = np.rint(u).astype(np.int64), np.rint(v).astype(np.int64)
    h, w = depth.shape
    check = valid & (ui >= 0) & (ui < w) & (vi >= 0) & (vi < h)
    keep = np.ones(len(pc), dtype=bool)
    idx = np.where(check)[0]
    zm = depth[vi[idx], ui[idx]]
    with np.errstate(invalid="ignore"):
        free = (zm > 0) & np.isfinite(zm) & (zc[idx] < (zm - np.float32(merge_radius)))
    keep[idx[free]] = False
    return keep


__all__ = [n for n in dir() if not n.startswith("_")]
_ = fm  # model import kept for callers that want radii
