"""Robot self-filter on the GPU — drop-in for ``TyGrit/kinematics/robot_filter.py:13-64``.

``filter_robot_points`` keeps the reference's signature and semantics (uniform ``sphere_radius``
for all spheres, the 4th column of the sphere table is ignored, strict ``d2 < r2`` removes, centres
built in fp64 from the given link poses and cast to f32).  The point/sphere test runs in
``tg_filter_points`` with the reference's exact f32 evaluation order, so the kept set is
bit-identical, not merely equal outside a tolerance band.

``filter_robot_points_batch`` is the additive batch surface: depth frames + robot states in, keep
masks out (``tg_self_filter_depth``), FK included.
"""

from __future__ import annotations

import numpy as np
import torch

from . import engine
from .fetch_model import fetch_spheres_dict


def sphere_centres_from_poses(link_poses: dict, spheres: dict) -> np.ndarray:
    """World centres (S,3) f32 exactly as the reference builds them (robot_filter.py:35-49)."""
    centres = []
    for link_name, local_spheres in spheres.items():
        if link_name not in link_poses:
            continue
        T = np.asarray(link_poses[link_name], dtype=np.float64)
        local = np.asarray(local_spheres, dtype=np.float64)[:, :3]
        centres.append(local @ T[:3, :3].T + T[:3, 3])
    if not centres:
        return np.empty((0, 3), dtype=np.float32)
    return np.vstack(centres).astype(np.float32)


def filter_robot_points(points, link_poses: dict, spheres: dict | None = None, sphere_radius: float = 0.08):
    """Remove points inside the robot's collision volume; returns the kept (M,3) points.

    ``points`` may be a numpy array (returns numpy, like the reference) or a CUDA tensor (returns a
    CUDA tensor, no host round trip).
    """
    spheres = fetch_spheres_dict() if spheres is None else spheres
    is_tensor = isinstance(points, torch.Tensor)
    if points.shape[0] == 0:
        return points
    centres = sphere_centres_from_poses(link_poses, spheres)
    if centres.shape[0] == 0:
        return points
    dev = points.device if is_tensor and points.is_cuda else torch.device("cuda", torch.cuda.current_device())
    p = points if is_tensor else torch.from_numpy(np.ascontiguousarray(points, dtype=np.float32))
    p = p.to(dev, torch.float32)
    keep = engine.filter_points_mask(p, torch.from_numpy(centres).to(dev), sphere_radius).bool()
    if is_tensor:
        return points[keep.to(points.device)]
    return points[keep.cpu().numpy()]


def filter_robot_points_batch(depth, intrinsics, states, sphere_radius: float = 0.08, z_range=(0.2, 3.0),
                              stride: int = 1, want_world: bool = False):
    """Batched self-filter over depth frames.

    depth (S,H,W) f32 CUDA, intrinsics (3,3), states (S,13) ``[x, y, theta, q0..q7, head_pan, head_tilt]``.
    Returns keep mask (S,H',W') uint8 (1 = valid depth in range and not on the robot)
    [and world points (S,H',W',3)].
    """
    return engine.self_filter_depth(depth, intrinsics, states, sphere_radius, z_range, stride, want_world)
