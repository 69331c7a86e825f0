"""``B200Scene`` — device-resident point-cloud world model satisfying TyGrit's ``Scene`` protocol.

Mirror of ``TyGrit/belief_state/pointcloud_scene.py:36-268`` (same constructor arguments, the same
three clouds S / M / G, the same five update modes and the same results row for row), with every
cloud kept as a CUDA tensor: a depth frame goes depth -> world cloud -> crop -> ground filter ->
robot self-filter -> mode-specific merge without leaving the GPU, and
``B200Planner.update_environment`` can take ``get_pointcloud_device()`` directly — the reference's
per-iteration ``pts.tolist()`` hand-off (``planning/fetch/vamp_preview.py:222-224``) disappears.

``get_pointcloud()`` returns a fresh numpy array as the protocol demands
(``belief_state/scene.py:17-48``).
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Callable

import numpy as np
import torch

from . import ingest
from .types import SensorSnapshot

_VALID_MODES = frozenset({"static", "latest", "accumulated", "combine", "ray_casting"})


@dataclass(frozen=True)
class SceneConfig:
    """``TyGrit/belief_state/config.py:9-18``."""

    ground_z_threshold: float = 0.3
    depth_range: tuple[float, float] = (0.2, 3.0)
    enable_ground_filter: bool = True
    merge_radius: float = 0.03
    downsample_voxel_size: float = 0.05
    crop_radius: float = 2.5


@dataclass(frozen=True)
class PointCloudSceneConfig(SceneConfig):
    """``TyGrit/belief_state/config.py:21-26``."""

    mode: str = "latest"
    depth_stride: int = 4


# (points_world, snapshot) -> filtered points; numpy in/out like the reference's RobotFilterFn
RobotFilterFn = Callable[[np.ndarray, SensorSnapshot], np.ndarray]


def fetch_robot_centres(snapshot, device) -> torch.Tensor:
    """(111,3) f32 CUDA sphere centres of the Fetch in ``snapshot.robot_state`` (10-DOF FK incl. head)."""
    from . import engine
    from .types import base_xyt

    rs = snapshot.robot_state
    arm = torch.tensor([list(rs.planning_joints)], dtype=torch.float32, device=device)
    head = torch.tensor([list(rs.head_joints)], dtype=torch.float32, device=device)
    base = torch.tensor([list(base_xyt(rs.base_pose))], dtype=torch.float32, device=device)
    return engine.fk_spheres(arm, base, head)[0]


class B200Scene:
    """Point-cloud world model with the reference's five update modes, resident on one GPU.

    ``robot_filter`` selects the self-filter: ``None`` (off), ``"fetch"`` (built-in: FK of the
    snapshot's robot state + ``tg_ingest_depth``'s fused sphere test, radius ``robot_radius``), or a
    reference-style callable ``(points numpy, snapshot) -> points numpy`` (host round trip, kept for
    drop-in compatibility with ``PointCloudScene(robot_filter_fn=...)``).
    """

    def __init__(self, config: PointCloudSceneConfig | None = None, static_map=None, *, robot_filter_fn=None,
                 robot_filter: str | None = None, robot_radius: float = 0.08, device: int | str | torch.device = 0) -> None:
        cfg = config or PointCloudSceneConfig()
        if cfg.mode not in _VALID_MODES:
            raise ValueError(f"mode must be one of {sorted(_VALID_MODES)}, got {cfg.mode!r}")
        if robot_filter not in (None, "fetch"):
            raise ValueError(f"robot_filter must be None or 'fetch', got {robot_filter!r}")
        if not torch.cuda.is_available():
            raise RuntimeError("tygrit_b200: no CUDA device — B200Scene has no CPU fallback")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        raw_s = self._to_dev(static_map) if static_map is not None else self._empty()
        self._S = ingest.voxel_downsample(raw_s, cfg.downsample_voxel_size) if raw_s.shape[0] > 0 else raw_s
        self._M = self._empty()
        self._G = self._empty()
        self.mode = cfg.mode
        self.downsample_voxel_size = cfg.downsample_voxel_size
        self.merge_radius = cfg.merge_radius
        self.ground_z_threshold = cfg.ground_z_threshold
        self.enable_ground_filter = cfg.enable_ground_filter
        self.depth_range = cfg.depth_range
        self.crop_radius = cfg.crop_radius
        self._robot_filter_fn = robot_filter_fn
        self._robot_filter = robot_filter
        self._robot_radius = robot_radius
        self._depth_stride = cfg.depth_stride

    # ── helpers ──────────────────────────────────────────────────────────
    def _empty(self) -> torch.Tensor:
        return torch.empty(0, 3, dtype=torch.float32, device=self.device)

    def _to_dev(self, a) -> torch.Tensor:
        if isinstance(a, torch.Tensor):
            return a.to(self.device, torch.float32).reshape(-1, 3).contiguous()
        return torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(-1, 3))).to(self.device)

    # ── Scene protocol ───────────────────────────────────────────────────
    def update(self, snapshot: SensorSnapshot, camera_pose) -> None:
        """Integrate a new sensor observation (pointcloud_scene.py:92-135)."""
        if self.mode == "static":
            return
        depth = snapshot.depth
        d = depth if isinstance(depth, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(depth, dtype=np.float32))
        d = d.to(self.device, torch.float32)
        T_wc = np.asarray(camera_pose, dtype=np.float64).reshape(4, 4)
        K = np.asarray(snapshot.intrinsics, dtype=np.float64).reshape(3, 3)
        centres = fetch_robot_centres(snapshot, self.device) if self._robot_filter == "fetch" else None
        # steps 1-4 fused: back-projection, crop around the camera, ground filter, robot self-filter
        pcd = ingest.depth_to_world_pointcloud(
            d, K, T_wc, z_range=self.depth_range, stride=self._depth_stride, crop_radius=self.crop_radius,
            ground_z=self.ground_z_threshold if self.enable_ground_filter else None,
            robot_centres=centres, robot_radius=self._robot_radius)
        if self._robot_filter_fn is not None and pcd.shape[0] > 0:
            pcd = self._to_dev(self._robot_filter_fn(pcd.cpu().numpy(), snapshot))
        # step 5: mode-specific integration
        if self.mode == "latest":
            self._M = pcd
        elif self.mode == "accumulated":
            self._update_accumulated(pcd, K, T_wc)
        elif self.mode == "combine":
            self._update_combine(pcd)
        elif self.mode == "ray_casting":
            self._update_ray_casting(pcd, d, K, T_wc)

    def get_pointcloud(self) -> np.ndarray:
        """E(t) = S ∪ M(t) as a fresh (N,3) f32 numpy array."""
        return self._current_environment().cpu().numpy()

    def get_pointcloud_device(self) -> torch.Tensor:
        """E(t) as a CUDA tensor (what ``B200Planner.update_environment`` consumes without a copy)."""
        return self._current_environment()

    def clear(self) -> None:
        self._M = self._empty()

    # ── extra API (beyond protocol) ──────────────────────────────────────
    def set_goal_pcd(self, pcd_world) -> None:
        self._G = self._to_dev(pcd_world)

    def get_goal_pcd(self) -> np.ndarray:
        return self._G.cpu().numpy()

    def current_observations(self) -> np.ndarray:
        return self._M.cpu().numpy()

    # ── internals ────────────────────────────────────────────────────────
    def _current_environment(self, roi_center=None, roi_radius=None) -> torch.Tensor:
        if self.mode == "static":
            env = self._S.clone()
        elif self._S.shape[0] == 0:
            env = self._M.clone()
        elif self._M.shape[0] == 0:
            env = self._S.clone()
        else:
            env = ingest.merge_dedup(self._S, self._M, self.merge_radius)
        if roi_center is not None and roi_radius is not None and env.shape[0] > 0:
            env = ingest.crop_sphere(env, np.asarray(roi_center, dtype=np.float32), roi_radius)
        return env

    def _update_accumulated(self, pcd, K, T_wc) -> None:
        if pcd.shape[0] == 0:
            return
        M = self._M
        if M.shape[0] > 0:
            keep = ingest.carve_flags(M, "frustum", K, T_wc, self.depth_range, 640, 480)
            M = ingest.compact(M, keep)
        self._M = ingest.merge_dedup(M, pcd, self.merge_radius)

    def _update_combine(self, pcd) -> None:
        if pcd.shape[0] == 0:
            return
        M = pcd if self._M.shape[0] == 0 else torch.cat((self._M, pcd), dim=0)
        self._M = ingest.voxel_downsample(M, self.downsample_voxel_size)

    def _update_ray_casting(self, pcd, depth, K, T_wc) -> None:
        M = self._M
        if M.shape[0] > 0:
            keep = ingest.carve_flags(M, "ray_casting", K, T_wc, self.depth_range, 0, 0, depth=depth,
                                      merge_radius=self.merge_radius)
            M = ingest.compact(M, keep)
        if pcd.shape[0] > 0:
            M = ingest.merge_dedup(M, pcd, self.merge_radius)
        self._M = M


__all__ = ["B200Scene", "PointCloudSceneConfig", "SceneConfig", "fetch_robot_centres"]
