"""Device-resident point-cloud utilities: same names and results as ``TyGrit/utils/pointcloud.py`` and
``TyGrit/utils/depth.py``, CUDA tensors in and out (no host round trip, no ``tolist()``).

Every function is a per-point predicate kernel followed by one order-preserving compaction, so the
output equals the reference's numpy result row for row (``tests/test_gpu_ingest.py``).  The only
synchronisation is the read-back of the kept-point count that sizes the returned tensor.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as N


def _stream(dev: torch.device) -> int:
    return int(torch.cuda.current_stream(dev).cuda_stream)


def _pts(t, name: str = "points") -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(np.asarray(t, dtype=np.float32))
    if not t.is_cuda:
        t = t.cuda()
    t = t.to(torch.float32).contiguous()
    if t.ndim != 2 or t.shape[1] != 3:
        raise ValueError(f"{name}: expected shape (N, 3), got {tuple(t.shape)}")
    return t


def _f64(a, shape) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))


def compact(points: torch.Tensor, flags: torch.Tensor) -> torch.Tensor:
    """``points[flags != 0]`` in input order (numpy boolean-mask order)."""
    p = _pts(points)
    n = p.shape[0]
    if n == 0:
        return p
    f = flags.to(device=p.device, dtype=torch.uint8).contiguous()
    out = torch.empty_like(p)
    cnt = torch.zeros(1, dtype=torch.int32, device=p.device)
    with torch.cuda.device(p.device):
        N.check(N.lib().tg_compact_points(p.data_ptr(), f.data_ptr(), n, out.data_ptr(), n, cnt.data_ptr(), _stream(p.device)),
                "tg_compact_points")
    return out[: int(cnt.item())]


def voxel_first_flags(base: torch.Tensor | None, add: torch.Tensor, voxel: float) -> torch.Tensor:
    """uint8 flags over ``add``: first point of every voxel that no ``base`` point occupies."""
    a = _pts(add, "add")
    b = _pts(base, "base") if base is not None and len(base) else None
    keep = torch.empty(a.shape[0], dtype=torch.uint8, device=a.device)
    if a.shape[0] == 0:
        return keep
    with torch.cuda.device(a.device):
        N.check(
            N.lib().tg_voxel_first(b.data_ptr() if b is not None else None, b.shape[0] if b is not None else 0, a.data_ptr(),
                                   a.shape[0], float(voxel), keep.data_ptr(), None, _stream(a.device)),
            "tg_voxel_first",
        )
    return keep


def voxel_downsample(points: torch.Tensor, voxel_size: float) -> torch.Tensor:
    """``utils/pointcloud.py:13-31``: the first point of every voxel, input order kept."""
    p = _pts(points)
    if p.shape[0] == 0 or voxel_size <= 0:
        return p
    return compact(p, voxel_first_flags(None, p, max(voxel_size, 1e-9)))


def merge_dedup(base: torch.Tensor, add: torch.Tensor, radius: float) -> torch.Tensor:
    """``utils/pointcloud.py:34-64``: all of ``base`` plus the first new-voxel points of ``add``."""
    b, a = _pts(base, "base"), _pts(add, "add")
    if a.shape[0] == 0:
        return b
    if b.shape[0] == 0:
        return a
    kept = compact(a, voxel_first_flags(b, a, max(radius, 1e-6)))
    return b if kept.shape[0] == 0 else torch.cat((b, kept), dim=0)


def crop_sphere(points: torch.Tensor, center, radius: float) -> torch.Tensor:
    """``utils/pointcloud.py:67-80``: XY cylinder, inclusive, f32."""
    p = _pts(points)
    if p.shape[0] == 0:
        return p
    c = torch.as_tensor(np.asarray(center, dtype=np.float32)[:2], device=p.device)
    d = p[:, :2] - c
    d2 = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]
    return compact(p, (d2 <= np.float32(radius * radius)).to(torch.uint8))


def filter_ground(points: torch.Tensor, z_threshold: float) -> torch.Tensor:
    """``utils/pointcloud.py:83-90``: keep z > threshold."""
    p = _pts(points)
    if p.shape[0] == 0:
        return p
    return compact(p, (p[:, 2] > np.float32(z_threshold)).to(torch.uint8))


def depth_to_world_pointcloud(depth: torch.Tensor, intrinsics, extrinsics, z_range=(0.2, 3.0), stride: int = 4, *,
                              crop_radius: float | None = None, ground_z: float | None = None,
                              robot_centres: torch.Tensor | None = None, robot_radius: float = 0.08) -> torch.Tensor:
    """``utils/depth.py:52-84`` on the device, optionally fused with the crop / ground / robot filters of
    ``PointCloudScene.update`` (``belief_state/pointcloud_scene.py:100-121``).  Returns (N,3) f32 CUDA."""
    d = depth if isinstance(depth, torch.Tensor) else torch.as_tensor(np.asarray(depth, dtype=np.float32))
    if not d.is_cuda:
        d = d.cuda()
    d = d.to(torch.float32).contiguous()
    if d.ndim != 2:
        raise ValueError("depth must be (H, W)")
    H, W = d.shape
    K, T = _f64(intrinsics, (3, 3)), _f64(extrinsics, (4, 4))
    cap = ((H + stride - 1) // stride) * ((W + stride - 1) // stride)
    out = torch.empty(cap, 3, dtype=torch.float32, device=d.device)
    cnt = torch.zeros(1, dtype=torch.int32, device=d.device)
    cen = None
    if robot_centres is not None and len(robot_centres):
        cen = _pts(robot_centres, "robot_centres")
    with torch.cuda.device(d.device):
        N.check(
            N.lib().tg_ingest_depth(d.data_ptr(), H, W, int(stride), K.ctypes.data, T.ctypes.data, float(z_range[0]),
                                    float(z_range[1]), float(crop_radius) if crop_radius is not None else -1.0,
                                    float(ground_z) if ground_z is not None else float("nan"),
                                    cen.data_ptr() if cen is not None else None, cen.shape[0] if cen is not None else 0,
                                    float(robot_radius), out.data_ptr(), cap, cnt.data_ptr(), _stream(d.device)),
            "tg_ingest_depth",
        )
    return out[: int(cnt.item())]


def carve_flags(points: torch.Tensor, mode: str, intrinsics, extrinsics, z_range, width: int, height: int,
                depth: torch.Tensor | None = None, merge_radius: float = 0.03) -> torch.Tensor:
    """keep-flags of the accumulated map: ``mode="frustum"`` (``utils/pointcloud.py:93-133``, 0 = inside the
    frustum) or ``mode="ray_casting"`` (``pointcloud_scene.py:218-262``, 0 = seen through)."""
    p = _pts(points)
    keep = torch.empty(p.shape[0], dtype=torch.uint8, device=p.device)
    if p.shape[0] == 0:
        return keep
    K, T = _f64(intrinsics, (3, 3)), _f64(extrinsics, (4, 4))
    m = {"frustum": 0, "ray_casting": 1}[mode]
    dd = None
    if m == 1:
        dd = depth.to(device=p.device, dtype=torch.float32).contiguous()
        height, width = dd.shape
    with torch.cuda.device(p.device):
        N.check(
            N.lib().tg_carve_flags(p.data_ptr(), p.shape[0], m, K.ctypes.data, T.ctypes.data, float(z_range[0]), float(z_range[1]),
                                   int(width), int(height), dd.data_ptr() if dd is not None else None, float(merge_radius),
                                   keep.data_ptr(), _stream(p.device)),
            "tg_carve_flags",
        )
    return keep


__all__ = ["compact", "voxel_first_flags", "voxel_downsample", "merge_dedup", "crop_sphere", "filter_ground",
           "depth_to_world_pointcloud", "carve_flags"]
_ = C
