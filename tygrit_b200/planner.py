"""``B200Planner`` — drop-in for ``VampPreviewPlanner`` behind the ``MotionPlanner`` protocol.

Surface mirrored (names, argument meaning, error behaviour):
``TyGrit/planning/motion_planner.py:15-64`` (protocol) and the concrete helpers of
``TyGrit/planning/fetch/vamp_preview.py:72-338``.  Expected failures are returned as
``PlanResult(success=False, failure=...)``, never raised (``docs/architecture.md:19``).

Every verdict comes from the CUDA engine (``CollisionWorld``): single configurations through the
same kernels as batches.  On top of the reference surface there are tensor/array batch
extensions (``validate_configs``, ``validate_edges``) — one Python call per configuration cannot
reach the throughput the kernels have.

The planners themselves (VAMP's RRT-Connect / multilayer RRTC live in the absent third-party
library, ``planning/fetch/algorithms.py:23-147``) are re-designed for batch edge validation:
every iteration proposes ``rrt_batch`` extensions, validates them in one ``tg_check_edges`` call,
and tries to connect the trees with a second batched call.
"""

from __future__ import annotations

import time
from typing import Any, Sequence

import numpy as np

from .config import B200PlannerConfig, PlannerConfig
from .engine import CollisionWorld
from .fetch_model import JOINT_LOWER, JOINT_UPPER
from .types import PlannerFailure, PlanResult, SE2Pose, WholeBodyConfig, base_xyt, make_trajectory

try:  # the reference logs through loguru; stay quiet without it
    from loguru import logger as log
except Exception:  # pragma: no cover
    import logging

    log = logging.getLogger("tygrit_b200")


def _halton(n: int, dim: int, skip: int) -> np.ndarray:
    """(n, dim) Halton points (the reference seeds VAMP's halton sampler, vamp_preview.py:58-59)."""
    primes = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)[:dim]
    idx = np.arange(skip + 1, skip + n + 1, dtype=np.int64)
    out = np.empty((n, dim))
    for d, b in enumerate(primes):
        f, r, i = 1.0, np.zeros(n), idx.copy()
        while i.max() > 0:
            f /= b
            r += f * (i % b)
            i //= b
        out[:, d] = r
    return out


class B200Planner:
    """CUDA collision engine behind TyGrit's ``MotionPlanner`` surface."""

    def __init__(self, config: B200PlannerConfig | None = None) -> None:
        self._config = config or B200PlannerConfig()
        cfg = self._config
        self._world = CollisionWorld(
            device=cfg.device, max_points=cfg.max_points, voxels_per_metre=cfg.voxels_per_metre,
            point_radius=cfg.point_radius,
        )
        self._base = (0.0, 0.0, 0.0)          # module-level base pose used by arm-only planning
        self._obstacles: list[list[float]] = []
        self._attached = False
        self._pc_bounds_xy: tuple[float, float, float, float] | None = None
        self._halton_skip = 0
        log.debug("B200Planner initialised (algorithm={})", cfg.algorithm)

    @property
    def world(self) -> CollisionWorld:
        return self._world

    # ------------------------------------------------------------------ #
    # MotionPlanner protocol
    # ------------------------------------------------------------------ #
    def update_environment(self, points, base_pose) -> None:
        """Sync collision world: set base pose and replace the point cloud (vamp_preview.py:183-192)."""
        self.set_base_params(base_pose)
        if hasattr(points, "is_cuda") and points.is_cuda:   # device-resident cloud: no host round trip
            if points.numel() > 0:
                self.add_pointcloud(points)
            else:
                self.clear_pointclouds()
            return
        pts = np.asarray(points)
        if pts.size > 0:
            self.add_pointcloud(pts)
        else:
            self.clear_pointclouds()

    def validate_config(self, config: WholeBodyConfig) -> bool:
        """True iff the configuration is collision-free (vamp_preview.py:159-163)."""
        arm = np.asarray(config.arm_joints, dtype=np.float32).reshape(1, 8)
        base = np.asarray(base_xyt(config.base_pose), dtype=np.float32).reshape(1, 3)
        return bool(self._world.check_configs_host(arm, base)[0])

    def plan_interpolation(self, start, goal, base_pose) -> PlanResult:
        """Linear joint-space interpolation, no collision checking (vamp_preview.py:165-181)."""
        start = np.asarray(start, dtype=np.float64)
        goal = np.asarray(goal, dtype=np.float64)
        n = max(2, self._config.interpolation_steps)
        alphas = np.linspace(0.0, 1.0, n)
        arm_path = [(start + a * (goal - start)).tolist() for a in alphas]
        bx, by, bt = base_xyt(base_pose)
        return PlanResult(success=True, trajectory=make_trajectory(arm_path, [[bx, by, bt]] * n))

    def plan_arm(self, start, goal) -> PlanResult:
        """Arm-only plan between two 8-DOF configurations at the current base (vamp_preview.py:72-92)."""
        t0 = time.perf_counter()
        start = np.asarray(start, dtype=np.float64).reshape(8)
        goal = np.asarray(goal, dtype=np.float64).reshape(8)
        b = np.asarray(self._base, dtype=np.float64)
        lo = np.concatenate([b, JOINT_LOWER])
        hi = np.concatenate([b, JOINT_UPPER])
        path, iters = self._rrt_connect(np.concatenate([b, start]), np.concatenate([b, goal]), lo, hi)
        stats = {"planning_time_ms": (time.perf_counter() - t0) * 1e3, "iterations": float(iters)}
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
        if self._config.simplify:
            path = self._shortcut(path)
        path = self._interpolate_steps(path, self._config.interpolation_steps)
        # arm-only trajectories carry base [0, 0, 0] for every waypoint (algorithms.py:56-58)
        return PlanResult(success=True, trajectory=make_trajectory(path[:, 3:].tolist(), [[0.0, 0.0, 0.0]] * len(path)),
                          stats=stats)

    def plan_whole_body(self, start: WholeBodyConfig, goal: WholeBodyConfig) -> PlanResult:
        """Arm + base plan; start / goal collisions are reported as failures (vamp_preview.py:94-157)."""
        t0 = time.perf_counter()
        s = np.concatenate([base_xyt(start.base_pose), np.asarray(start.arm_joints, dtype=np.float64)])
        g = np.concatenate([base_xyt(goal.base_pose), np.asarray(goal.arm_joints, dtype=np.float64)])
        if self._is_in_collision(s[3:], s[:3]):
            return PlanResult(success=False, failure=PlannerFailure.COLLISION_AT_START)
        if self._is_in_collision(g[3:], g[:3]):
            return PlanResult(success=False, failure=PlannerFailure.COLLISION_AT_GOAL)
        pad = self._config.bounds_padding
        if self._pc_bounds_xy is not None:
            x0, x1, y0, y1 = self._pc_bounds_xy
        else:
            x0, x1 = min(s[0], g[0]) - 1.0, max(s[0], g[0]) + 1.0
            y0, y1 = min(s[1], g[1]) - 1.0, max(s[1], g[1]) + 1.0
        x0, x1 = min(x0, s[0], g[0]) - pad, max(x1, s[0], g[0]) + pad
        y0, y1 = min(y0, s[1], g[1]) - pad, max(y1, s[1], g[1]) + pad
        lo = np.concatenate([[x0, y0, -np.pi], JOINT_LOWER])
        hi = np.concatenate([[x1, y1, np.pi], JOINT_UPPER])
        path, iters = self._rrt_connect(s, g, lo, hi)
        stats = {"planning_time_ms": (time.perf_counter() - t0) * 1e3, "iterations": float(iters)}
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
        if self._config.simplify:
            path = self._shortcut(path)
        path = self._interpolate_density(path, self._config.interpolation_density)
        return PlanResult(success=True, trajectory=make_trajectory(path[:, 3:].tolist(), path[:, :3].tolist()), stats=stats)

    # ------------------------------------------------------------------ #
    # environment management (concrete-only helpers, vamp_preview.py:198-322)
    # ------------------------------------------------------------------ #
    def set_base_params(self, base_pose) -> None:
        self._base = base_xyt(base_pose)

    def add_pointcloud(self, points, point_radius: float | None = None) -> None:
        """Replace the collision cloud; every point is a sphere of ``point_radius`` (vamp_preview.py:205-236)."""
        radius = point_radius if point_radius is not None else self._config.point_radius
        self._world.set_point_radius(float(radius))
        is_cuda = hasattr(points, "is_cuda") and points.is_cuda
        self._world.build(points if is_cuda else np.asarray(points, dtype=np.float32))
        pad = self._config.bounds_padding
        if is_cuda:  # device cloud (B200Scene.get_pointcloud_device): only the four XY bounds come back
            if points.numel() > 0 and points.ndim == 2 and points.shape[1] >= 2:
                lo, hi = points[:, :2].amin(dim=0), points[:, :2].amax(dim=0)
                import torch

                b = torch.cat((lo, hi)).tolist()
                self._pc_bounds_xy = (b[0] - pad, b[2] + pad, b[1] - pad, b[3] + pad)
            return
        pts = np.asarray(points)
        if pts.size > 0 and pts.ndim == 2 and pts.shape[1] >= 2:
            mn, mx = pts[:, :2].min(axis=0), pts[:, :2].max(axis=0)
            self._pc_bounds_xy = (float(mn[0] - pad), float(mx[0] + pad), float(mn[1] - pad), float(mx[1] + pad))

    def clear_pointclouds(self) -> None:
        self._world.build(np.empty((0, 3), dtype=np.float32))

    def add_sphere(self, position: Sequence[float], radius: float) -> None:
        self._obstacles.append([float(position[0]), float(position[1]), float(position[2]), float(radius)])
        self._world.set_obstacles(self._obstacles)

    def clear_spheres(self) -> None:
        self._obstacles = []
        self._world.set_obstacles(np.empty((0, 4), dtype=np.float32))

    def attach_to_eef(self, spheres: list[dict[str, Any]], offset_pos=None, offset_quat=None) -> bool:
        """Attach collision spheres to the end-effector; False if something is already attached."""
        if self._attached:
            log.warning("Cannot attach: object already attached — detach first.")
            return False
        xyzr = [[*map(float, s["position"]), float(s["radius"])] for s in spheres]
        self._world.set_attachment(xyzr, offset_pos, offset_quat)
        self._attached = True
        return True

    def detach_from_eef(self) -> bool:
        if not self._attached:
            log.warning("Cannot detach: nothing attached.")
            return False
        self._world.set_attachment(np.empty((0, 4), dtype=np.float32))
        self._attached = False
        return True

    def check_trajectory_collisions(self, arm_path, base_configs) -> bool:
        """True iff any waypoint or any segment between consecutive waypoints collides (:306-322)."""
        arm = np.asarray(arm_path, dtype=np.float32).reshape(-1, 8)
        base = np.asarray(base_configs, dtype=np.float32).reshape(-1, 3)
        if arm.shape[0] != base.shape[0]:
            raise ValueError("arm_path and base_configs must have the same length")
        if arm.shape[0] == 0:
            return False
        if not bool(self._world.check_configs_host(arm[:1], base[:1])[0]):
            return True
        if arm.shape[0] == 1:
            return False
        ok, _ = self._world.check_edges_host(arm[:-1], base[:-1], arm[1:], base[1:], self._config.rake,
                                             self._config.resolution)
        return not bool(ok.all())

    # ------------------------------------------------------------------ #
    # batch extensions (tensors or arrays in, same kind out)
    # ------------------------------------------------------------------ #
    def validate_configs(self, arm, base, clearance: bool = False):
        """(B,8), (B,3)[x,y,theta] -> valid (B,) [, clearance (B,)]; CUDA tensors stay on the device."""
        if hasattr(arm, "is_cuda") and arm.is_cuda:
            out = self._world.check_configs(arm, base, clearance=clearance)
            return (out[0].bool(), out[1]) if clearance else out.bool()
        out = self._world.check_configs_host(np.asarray(arm), np.asarray(base), clearance=clearance)
        return (out[0].astype(bool), out[1]) if clearance else out.astype(bool)

    def validate_edges(self, arm0, base0, arm1, base1):
        if hasattr(arm0, "is_cuda") and arm0.is_cuda:
            return self._world.check_edges(arm0, base0, arm1, base1, self._config.rake, self._config.resolution).bool()
        ok, _ = self._world.check_edges_host(arm0, base0, arm1, base1, self._config.rake, self._config.resolution)
        return ok.astype(bool)

    # ------------------------------------------------------------------ #
    # internals
    # ------------------------------------------------------------------ #
    def _is_in_collision(self, arm_config, base_config) -> bool:
        arm = np.asarray(arm_config, dtype=np.float32).reshape(1, 8)
        base = np.asarray(base_config, dtype=np.float32).reshape(1, 3)
        return not bool(self._world.check_configs_host(arm, base)[0])

    def _edges_ok(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """Validate segments a[i] -> b[i] of 11-vectors [x, y, theta, q0..q7]."""
        a32, b32 = a.astype(np.float32), b.astype(np.float32)
        ok, _ = self._world.check_edges_host(a32[:, 3:], a32[:, :3], b32[:, 3:], b32[:, :3], self._config.rake,
                                             self._config.resolution)
        return ok.astype(bool)

    def _rrt_connect(self, start: np.ndarray, goal: np.ndarray, lo: np.ndarray, hi: np.ndarray):
        """Batched RRT-Connect over 11-vectors; returns (path (W,11) | None, iterations)."""
        cfg = self._config
        if self._is_in_collision(start[3:], start[:3]) or self._is_in_collision(goal[3:], goal[:3]):
            return None, 0
        if self._edges_ok(start[None], goal[None])[0]:
            return np.stack([start, goal]), 0
        free = hi > lo
        trees = [{"q": [start], "parent": [-1]}, {"q": [goal], "parent": [-1]}]
        deadline = time.perf_counter() + cfg.timeout
        for it in range(1, cfg.rrt_max_iterations + 1):
            if time.perf_counter() > deadline:
                return None, it
            ta, tb = trees[(it - 1) % 2], trees[it % 2]
            u = _halton(cfg.rrt_batch, int(free.sum()), self._halton_skip)
            self._halton_skip += cfg.rrt_batch
            samples = np.tile(lo, (cfg.rrt_batch, 1))
            samples[:, free] = lo[free] + u * (hi[free] - lo[free])
            qa = np.asarray(ta["q"])
            d = np.linalg.norm(samples[:, None, :] - qa[None, :, :], axis=2)
            near = d.argmin(axis=1)
            dist = d[np.arange(len(near)), near]
            scale = np.minimum(1.0, cfg.rrt_range / np.maximum(dist, 1e-12))
            new = qa[near] + scale[:, None] * (samples - qa[near])
            ok = self._edges_ok(qa[near], new)
            if not ok.any():
                continue
            new, near = new[ok], near[ok]
            first_new = len(ta["q"])
            ta["q"].extend(list(new))
            ta["parent"].extend(list(near))
            # connect: each new node to its nearest node in the other tree
            qb = np.asarray(tb["q"])
            db = np.linalg.norm(new[:, None, :] - qb[None, :, :], axis=2)
            nb = db.argmin(axis=1)
            okc = self._edges_ok(new, qb[nb])
            if okc.any():
                j = int(np.flatnonzero(okc)[np.argmin(db[np.arange(len(nb)), nb][okc])])
                pa = self._trace(ta, first_new + j)
                pb = self._trace(tb, int(nb[j]))
                path = pa[::-1] + pb                     # root_a .. new, nearest_b .. root_b
                if (it - 1) % 2 == 1:                    # ta is the goal tree: flip
                    path = path[::-1]
                return np.asarray(path), it
        return None, cfg.rrt_max_iterations

    @staticmethod
    def _trace(tree, i: int) -> list[np.ndarray]:
        out = []
        while i >= 0:
            out.append(tree["q"][i])
            i = tree["parent"][i]
        return out

    def _shortcut(self, path: np.ndarray) -> np.ndarray:
        """Batched shortcutting: test all (i, j) skips at once, take the longest valid one, repeat."""
        for _ in range(self._config.shortcut_rounds):
            n = len(path)
            if n <= 2:
                break
            ii, jj = np.triu_indices(n, k=2)
            ok = self._edges_ok(path[ii], path[jj])
            if not ok.any():
                break
            k = int(np.argmax(np.where(ok, jj - ii, -1)))
            path = np.concatenate([path[: ii[k] + 1], path[jj[k]:]])
        return path

    @staticmethod
    def _interpolate_steps(path: np.ndarray, steps: int) -> np.ndarray:
        """Subdivide every segment into ``steps`` equal parts (assumption for VAMP's path.interpolate(n))."""
        steps = max(1, int(steps))
        out = [path[0]]
        for a, b in zip(path[:-1], path[1:]):
            for k in range(1, steps + 1):
                out.append(a + (k / steps) * (b - a))
        return np.asarray(out)

    @staticmethod
    def _interpolate_density(path: np.ndarray, density: float) -> np.ndarray:
        """Subdivide so consecutive waypoints are at most ``density`` apart (wb.interpolate(density))."""
        out = [path[0]]
        for a, b in zip(path[:-1], path[1:]):
            n = max(1, int(np.ceil(np.linalg.norm(b - a) / max(density, 1e-9))))
            for k in range(1, n + 1):
                out.append(a + (k / n) * (b - a))
        return np.asarray(out)


def create_fetch_planner(planner: str, config: PlannerConfig | None = None) -> B200Planner:
    """Factory branch a maintainer adds to ``TyGrit/planning/fetch/__init__.py:9-32``."""
    if planner == "b200":
        if config is None:
            config = B200PlannerConfig()
        if not isinstance(config, B200PlannerConfig):
            raise TypeError(f"Expected B200PlannerConfig for b200 planner, got {type(config).__name__}")
        return B200Planner(config)
    raise ValueError(f"Unknown Fetch planner: {planner!r}. Available: b200")


def create_planner(robot: str, planner: str, config: PlannerConfig | None = None) -> B200Planner:
    """Mirror of ``TyGrit/planning/planner.py:16-39``."""
    if robot == "fetch":
        return create_fetch_planner(planner, config)
    raise ValueError(f"Unknown robot: {robot!r}")


def create_collision_check(planner: B200Planner):
    """``CollisionCheckFn`` factory (``TyGrit/checker/collision.py:22-27``)."""
    return planner.validate_config
