"""``B200Planner`` — drop-in for ``VampPreviewPlanner`` behind the ``MotionPlanner`` protocol.

Surface mirrored (names, argument meaning, error behaviour):
``TyGrit/planning/motion_planner.py:15-64`` (protocol) and the concrete helpers of
``TyGrit/planning/fetch/vamp_preview.py:72-338``.  Expected failures are returned as
``PlanResult(success=False, failure=...)``, never raised (``docs/architecture.md:19``).

Every verdict comes from the CUDA engine (``CollisionWorld``): single configurations through the
same kernels as batches.  On top of the reference surface there are tensor/array batch
extensions (``validate_configs``, ``validate_edges``) — one Python call per configuration cannot
reach the throughput the kernels have.

The planners themselves (VAMP's RRT-Connect / multilayer RRTC / FCIT* live in the absent third-party
library, ``planning/fetch/algorithms.py:23-232``) are re-designed as device-resident batch planners
(``csrc/planner.cu``): trees, lattice, roadmap, candidate edges and verdicts stay in HBM; the host enqueues kernels
and polls a status word; the finished path is the only device->host copy of a plan.

* ``algorithm = "rrtc"``: batched RRT-Connect (``rrt_batch`` extensions per iteration, Halton sampler).
* ``whole_body_algorithm = "multilayer_rrtc"``: base layer = SE(2) lattice search with ``reverse_penalty = 50``
  (``algorithms.py:93-96``), arm layer = RRT-Connect at the goal base, then whole-body shortcutting; falls back to
  an 11-D RRT-Connect (theta on the shortest arc) when a layer fails.
* ``whole_body_algorithm = "fcit_wb"``: fully connected roadmap inside the cloud's XY bounds, EVERY edge validated in one
  batched call (10^5-10^6 edges in a few ms).
Unknown names raise ``ValueError`` (the reference would fail inside ``configure_robot_and_planner_with_kwargs``).

Self-collision: VAMP's ``validate_whole_body_config`` also tests the robot against itself; its pair list is not
available, so ``self_collision=True`` (the default here) uses the reference's own ``SELF_COLLISION_PAIRS``
(``rl/train.py:78-107``).  The bare ``CollisionWorld`` keeps the term off.
"""

from __future__ import annotations

import ctypes as C
import time
from typing import Any, Sequence

import numpy as np

from . import _native as N
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




_ARM_ALGORITHMS = ("rrtc",)
_WB_ALGORITHMS = ("multilayer_rrtc", "fcit_wb", "rrtc_wb")
_SAMPLERS = ("halton",)
_MAX_PATH = 256


def _wrap(a):
    """Angle(s) to (-pi, pi]."""
    return -((-np.asarray(a, dtype=np.float64) + np.pi) % (2.0 * np.pi) - np.pi)


def _f32(v, n):
    a = np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(n))
    return a


class B200Planner:
    """CUDA collision engine + device-resident planners behind TyGrit's ``MotionPlanner`` surface."""

    def __init__(self, config: B200PlannerConfig | None = None) -> None:
        self._config = config or B200PlannerConfig()
        cfg = self._config
        # unknown names fail loudly (the reference fails inside VAMP's configure_robot_and_planner_with_kwargs)
        if cfg.algorithm not in _ARM_ALGORITHMS:
            raise ValueError(f"Unknown arm planner algorithm: {cfg.algorithm!r}. Available: {', '.join(_ARM_ALGORITHMS)}")
        if cfg.whole_body_algorithm not in _WB_ALGORITHMS:
            raise ValueError(f"Unknown whole-body algorithm: {cfg.whole_body_algorithm!r}. Available: {', '.join(_WB_ALGORITHMS)}")
        if cfg.sampler not in _SAMPLERS:
            raise ValueError(f"Unknown sampler: {cfg.sampler!r}. Available: {', '.join(_SAMPLERS)}")
        self._world = CollisionWorld(
            device=cfg.device, max_points=cfg.max_points, voxels_per_metre=cfg.voxels_per_metre,
            point_radius=cfg.point_radius,
        )
        self._world.set_self_collision(cfg.self_collision)
        self._lib = N.lib()
        h = C.c_void_p()
        N.check(self._lib.tg_planner_create(C.byref(h), self._world._h, int(cfg.rrt_max_nodes), int(cfg.rrt_batch)), "tg_planner_create")
        self._plan = h
        self._base = (0.0, 0.0, 0.0)          # module-level base pose used by arm-only planning
        self._obstacles: list[list[float]] = []
        self._attached = False
        self._pc_bounds_xy: tuple[float, float, float, float] | None = None
        self._clouds: list = []               # add_pointcloud accumulates until clear_pointclouds (like VAMP's Environment)
        self._halton_skip = int(cfg.seed) * 1_000_003
        log.debug("B200Planner initialised (algorithm={})", cfg.algorithm)

    def __del__(self) -> None:  # pragma: no cover
        try:
            if getattr(self, "_plan", None):
                self._lib.tg_planner_destroy(self._plan)
                self._plan = None
        except Exception:
            pass

    @property
    def world(self) -> CollisionWorld:
        return self._world

    def _stream(self) -> int:
        import torch

        return int(torch.cuda.current_stream(self._world.device).cuda_stream)

    # ------------------------------------------------------------------ #
    # MotionPlanner protocol
    # ------------------------------------------------------------------ #
    def update_environment(self, points, base_pose) -> None:
        """Sync collision world: set base pose and replace the point cloud (vamp_preview.py:183-192)."""
        self.set_base_params(base_pose)
        self.clear_pointclouds()
        if hasattr(points, "is_cuda") and points.is_cuda:   # device-resident cloud: no host round trip
            if points.numel() > 0:
                self.add_pointcloud(points)
            return
        pts = np.asarray(points)
        if pts.size > 0:
            self.add_pointcloud(pts)

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
        """Arm-only plan between two 8-DOF configurations at the current base (vamp_preview.py:72-92,
        algorithms.py:23-66): RRT-Connect, simplify, interpolate(interpolation_steps)."""
        t0 = time.perf_counter()
        start = np.asarray(start, dtype=np.float64).reshape(8)
        goal = np.asarray(goal, dtype=np.float64).reshape(8)
        b = np.asarray(self._base, dtype=np.float64)
        lo = np.concatenate([b, JOINT_LOWER])
        hi = np.concatenate([b, JOINT_UPPER])
        path, iters = self._rrt_connect(np.concatenate([b, start]), np.concatenate([b, goal]), lo, hi)
        stats = {"planning_time_ms": (time.perf_counter() - t0) * 1e3, "planning_iterations": float(iters)}
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
        raw = path
        if self._config.simplify:
            path = self._shortcut(path)
        path = self._final(path, raw, lambda p: self._interpolate_steps(p, self._config.interpolation_steps), stats)
        stats["planning_time_ms"] = (time.perf_counter() - t0) * 1e3
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
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
        cfg = self._config
        stats: dict[str, float] = {}
        if cfg.whole_body_algorithm == "fcit_wb":
            if self._pc_bounds_xy is None:
                log.warning("FCIT* selected but XY bounds unavailable — call add_pointcloud first.")
                return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND)
            path = self._fcit(s, g, stats)
        elif cfg.whole_body_algorithm == "multilayer_rrtc":
            path = self._multilayer(s, g, stats)
        else:
            path = None
        if path is None and cfg.whole_body_algorithm != "fcit_wb":   # 11-D RRT-Connect: explicit choice or fallback
            lo, hi = self._wb_box(s, g)
            path, iters = self._rrt_connect(s, g, lo, hi)
            stats["planning_iterations"] = float(iters)
        stats["total_planning_time_ms"] = (time.perf_counter() - t0) * 1e3
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
        raw = path
        if cfg.simplify or cfg.whole_body_algorithm == "multilayer_rrtc":   # whole_body_simplify is unconditional (algorithms.py:124)
            t1 = time.perf_counter()
            path = self._shortcut(path, cfg.fcit_reverse_weight if cfg.whole_body_algorithm == "fcit_wb" else cfg.reverse_penalty)
            stats["simplification_time_ms"] = (time.perf_counter() - t1) * 1e3
        path = self._final(path, raw, lambda p: self._interpolate_density(p, cfg.interpolation_density), stats)
        stats["total_planning_time_ms"] = (time.perf_counter() - t0) * 1e3
        if path is None:
            return PlanResult(success=False, failure=PlannerFailure.NO_PATH_FOUND, stats=stats)
        return PlanResult(success=True, trajectory=make_trajectory(path[:, 3:].tolist(), path[:, :3].tolist()), stats=stats)

    # ------------------------------------------------------------------ #
    # environment management (concrete-only helpers, vamp_preview.py:198-322)
    # ------------------------------------------------------------------ #
    def set_base_params(self, base_pose) -> None:
        self._base = base_xyt(base_pose)

    def add_pointcloud(self, points, point_radius: float | None = None) -> None:
        """Add a point cloud as collision geometry; every point is a sphere of ``point_radius`` (vamp_preview.py:205-236).
        Clouds ACCUMULATE until :meth:`clear_pointclouds`, like ``vamp_preview.Environment.add_pointcloud``: the structure
        is rebuilt over the concatenation (the radius of the latest call applies to all of it)."""
        import torch

        radius = point_radius if point_radius is not None else self._config.point_radius
        self._world.set_point_radius(float(radius))
        dev = self._world.device
        if hasattr(points, "is_cuda") and points.is_cuda:
            t = points.to(device=dev, dtype=torch.float32).reshape(-1, 3)
        else:
            t = torch.from_numpy(np.ascontiguousarray(np.asarray(points, dtype=np.float32).reshape(-1, 3))).to(dev)
        if t.shape[0] == 0:
            return
        total = sum(int(c.shape[0]) for c in self._clouds) + int(t.shape[0])
        if total > self._world.max_points:
            raise RuntimeError(f"tygrit_b200: {total} accumulated cloud points exceed max_points={self._world.max_points}")
        self._clouds.append(t)
        cloud = self._clouds[0] if len(self._clouds) == 1 else torch.cat(self._clouds)
        self._world.build(cloud.contiguous())
        pad = self._config.bounds_padding
        lo, hi = cloud[:, :2].amin(dim=0), cloud[:, :2].amax(dim=0)   # only the four XY bounds come back to the host
        b = torch.cat((lo, hi)).tolist()
        self._pc_bounds_xy = (b[0] - pad, b[2] + pad, b[1] - pad, b[3] + pad)

    def clear_pointclouds(self) -> None:
        self._clouds = []
        self._world.build(np.empty((0, 3), dtype=np.float32))

    def add_sphere(self, position: Sequence[float], radius: float) -> None:
        sphere = [float(position[0]), float(position[1]), float(position[2]), float(radius)]
        if len(self._obstacles) + 1 > N.TG_MAX_OBSTACLES:
            raise RuntimeError(f"tygrit_b200: at most {N.TG_MAX_OBSTACLES} sphere obstacles")   # nothing was changed
        self._world.set_obstacles(self._obstacles + [sphere])
        self._obstacles.append(sphere)

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
    # internals: thin callers of csrc/planner.cu
    # ------------------------------------------------------------------ #
    def _is_in_collision(self, arm_config, base_config) -> bool:
        arm = np.asarray(arm_config, dtype=np.float32).reshape(1, 8)
        base = np.asarray(base_config, dtype=np.float32).reshape(1, 3)
        return not bool(self._world.check_configs_host(arm, base)[0])

    def _wb_box(self, s: np.ndarray, g: np.ndarray):
        """Sampling box of whole-body planning: cloud XY bounds (or a box around start / goal), theta, joint limits."""
        pad = self._config.bounds_padding
        if self._pc_bounds_xy is not None:
            x0, x1, y0, y1 = self._pc_bounds_xy
        else:
            x0, x1 = min(s[0], g[0]) - 1.0, max(s[0], g[0]) + 1.0
            y0, y1 = min(s[1], g[1]) - 1.0, max(s[1], g[1]) + 1.0
        x0, x1 = min(x0, s[0], g[0]) - pad, max(x1, s[0], g[0]) + pad
        y0, y1 = min(y0, s[1], g[1]) - pad, max(y1, s[1], g[1]) + pad
        return np.concatenate([[x0, y0, -np.pi], JOINT_LOWER]), np.concatenate([[x1, y1, np.pi], JOINT_UPPER])

    def _rrt_connect(self, start: np.ndarray, goal: np.ndarray, lo: np.ndarray, hi: np.ndarray):
        """Device-resident batched RRT-Connect over 11-vectors; returns (path (W,11) f64 | None, iterations)."""
        cfg = self._config
        if self._is_in_collision(start[3:], start[:3]) or self._is_in_collision(goal[3:], goal[:3]):
            return None, 0
        out = np.empty((_MAX_PATH, 11), dtype=np.float32)
        n, iters = C.c_int(0), C.c_int(0)
        s32, g32, lo32, hi32 = _f32(start, 11), _f32(goal, 11), _f32(lo, 11), _f32(hi, 11)
        N.check(self._lib.tg_planner_rrt_connect(self._plan, s32.ctypes.data, g32.ctypes.data, lo32.ctypes.data, hi32.ctypes.data,
                                                 float(cfg.rrt_range), int(cfg.rrt_max_iterations), self._halton_skip & 0x7FFFFFFF,
                                                 int(cfg.rake), float(cfg.resolution), float(cfg.timeout), out.ctypes.data, _MAX_PATH,
                                                 C.byref(n), C.byref(iters), self._stream()), "tg_planner_rrt_connect")
        self._halton_skip += max(1, iters.value) * cfg.rrt_batch
        if n.value <= 0:
            return None, iters.value
        path = out[: n.value].astype(np.float64)
        path[0], path[-1, :2], path[-1, 3:] = start, goal[:2], goal[3:]      # exact endpoints (the trees hold f32 copies)
        return path, iters.value

    def _shortcut(self, path: np.ndarray, reverse_weight: float = 1.0) -> np.ndarray:
        """All waypoint pairs validated in one batch, then the shortest path through the valid skips (edges that drive the
        base backwards cost ``reverse_weight`` x their length)."""
        out = path
        for _ in range(max(1, self._config.shortcut_rounds)):
            if len(out) <= 2:
                break
            if len(out) > 128:                       # thin out evenly first (rare: very long lattice paths)
                keep = np.unique(np.linspace(0, len(out) - 1, 128).round().astype(int))
                if not self._edges_ok(out[keep[:-1]], out[keep[1:]]).all():
                    break
                out = out[keep]
            buf = np.ascontiguousarray(out, dtype=np.float32)
            n = C.c_int(len(buf))
            N.check(self._lib.tg_planner_shortcut(self._plan, buf.ctypes.data, C.byref(n), int(self._config.rake),
                                                  float(self._config.resolution), float(reverse_weight), self._stream()),
                    "tg_planner_shortcut")
            if n.value >= len(out):
                break
            new = buf[: n.value].astype(np.float64)
            new[0], new[-1] = out[0], out[-1]
            out = new
        return out

    def _final(self, path: np.ndarray, raw: np.ndarray, interpolate, stats: dict):
        """Interpolate and re-validate the trajectory that will be returned.  Edge validation samples at a fixed
        resolution, so the interpolated waypoints' own (shorter, more densely sampled) segments can expose a grazing
        contact the long edge stepped over; a planner must not hand out a trajectory its own
        ``check_trajectory_collisions`` rejects.  Falls back to the un-simplified path, then reports failure."""
        for cand in (path, raw):
            out = interpolate(cand)
            if self._edges_ok(out[:-1], out[1:]).all():
                return out
            stats["final_check_retries"] = stats.get("final_check_retries", 0.0) + 1.0
            if cand is raw:
                break
        return None

    def _edges_ok(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """Validate segments a[i] -> b[i] of 11-vectors [x, y, theta, q0..q7] (theta of b unwrapped relative to a)."""
        a32, b32 = a.astype(np.float32), b.astype(np.float32)
        b32[:, 2] = a32[:, 2] + _wrap(b32[:, 2] - a32[:, 2]).astype(np.float32)
        ok, _ = self._world.check_edges_host(a32[:, 3:], a32[:, :3], b32[:, 3:], b32[:, :3], self._config.rake,
                                             self._config.resolution)
        return ok.astype(bool)

    def _multilayer(self, s: np.ndarray, g: np.ndarray, stats: dict) -> np.ndarray | None:
        """multilayer_rrtc (algorithms.py:77-147): base layer on the SE(2) lattice with the arm held at the start
        configuration (reverse_penalty = 50), then the arm layer (RRT-Connect) at the goal base."""
        cfg = self._config
        base_moves = np.hypot(g[0] - s[0], g[1] - s[1]) > 1e-6 or abs(_wrap(g[2] - s[2])) > 1e-6
        t0 = time.perf_counter()
        path = s[None, :].copy()
        if base_moves:
            lo, hi = self._wb_box(s, g)
            bounds = _f32([lo[0], hi[0], lo[1], hi[1]], 4)
            out = np.empty((2048, 3), dtype=np.float32)
            n = C.c_int(0)
            st3 = (C.c_double * 3)()
            s3, g3, a8 = _f32(s[:3], 3), _f32(g[:3], 3), _f32(s[3:], 8)    # named: the arrays must outlive the call
            N.check(self._lib.tg_planner_base_path(self._plan, s3.ctypes.data, g3.ctypes.data,
                                                   a8.ctypes.data, bounds.ctypes.data, float(cfg.base_resolution),
                                                   float(cfg.reverse_penalty), float(cfg.base_turn_cost), int(cfg.rake), float(cfg.resolution),
                                                   int(cfg.base_max_rounds), out.ctypes.data, 2048, C.byref(n), st3, self._stream()),
                    "tg_planner_base_path")
            stats["base_planning_time_ms"] = (time.perf_counter() - t0) * 1e3
            stats["base_relaxation_sweeps"], stats["base_validation_rounds"], stats["base_lattice_nodes"] = (float(v) for v in st3)
            if n.value < 2:
                return None
            wp = out[: n.value].astype(np.float64)
            wp[0, :2], wp[-1, :2] = s[:2], g[:2]
            # collinear lattice steps are merged (the edge between the run's ends is the union of its steps)
            keep = [0]
            for i in range(1, len(wp) - 1):
                d0, d1 = wp[i] - wp[keep[-1]], wp[i + 1] - wp[i]
                if abs(d0[2]) > 1e-6 or abs(d1[2]) > 1e-6 or abs(d0[0] * d1[1] - d0[1] * d1[0]) > 1e-9 or (d0[:2] @ d1[:2]) <= 0:
                    keep.append(i)
            keep.append(len(wp) - 1)
            path = np.concatenate([wp, np.tile(s[3:], (len(wp), 1))], axis=1)
            merged = path[keep]
            if len(merged) < len(path) and self._edges_ok(merged[:-1], merged[1:]).all():
                path = merged
        t1 = time.perf_counter()
        if np.abs(g[3:] - s[3:]).max() > 1e-9:
            gb = path[-1, :3]                                   # goal base pose, theta continuous with the base path
            lo = np.concatenate([gb, JOINT_LOWER])
            hi = np.concatenate([gb, JOINT_UPPER])
            arm_path, iters = self._rrt_connect(np.concatenate([gb, s[3:]]), np.concatenate([gb, g[3:]]), lo, hi)
            stats["arm_planning_time_ms"] = (time.perf_counter() - t1) * 1e3
            stats["planning_iterations"] = float(iters)
            if arm_path is None:
                return None
            path = np.concatenate([path, arm_path[1:]])
        return path

    def _fcit(self, s: np.ndarray, g: np.ndarray, stats: dict) -> np.ndarray | None:
        """fcit_wb (algorithms.py:155-232): fully connected, fully validated roadmap inside the cloud's XY bounds."""
        cfg = self._config
        x0, x1, y0, y1 = self._pc_bounds_xy
        x0, x1 = min(x0, s[0], g[0]), max(x1, s[0], g[0])
        y0, y1 = min(y0, s[1], g[1]), max(y1, s[1], g[1])
        lo = np.concatenate([[x0, y0, -np.pi], JOINT_LOWER])
        hi = np.concatenate([[x1, y1, np.pi], JOINT_UPPER])
        out = np.empty((_MAX_PATH, 11), dtype=np.float32)
        n = C.c_int(0)
        st4 = (C.c_double * 4)()
        t0 = time.perf_counter()
        s32, g32, lo32, hi32 = _f32(s, 11), _f32(g, 11), _f32(lo, 11), _f32(hi, 11)   # named: the arrays must outlive the call
        N.check(self._lib.tg_planner_fcit(self._plan, s32.ctypes.data, g32.ctypes.data, lo32.ctypes.data,
                                          hi32.ctypes.data, int(cfg.fcit_batch_size), int(max(cfg.fcit_max_samples, cfg.fcit_batch_size)),
                                          int(cfg.fcit_max_iterations), float(cfg.fcit_reverse_weight), float(cfg.fcit_max_edge),
                                          self._halton_skip & 0x7FFFFFFF, int(cfg.rake), float(cfg.resolution), float(cfg.timeout),
                                          out.ctypes.data, _MAX_PATH, C.byref(n), st4, self._stream()), "tg_planner_fcit")
        stats["arm_planning_time_ms"] = (time.perf_counter() - t0) * 1e3
        stats["fcit_samples"], stats["fcit_nodes"], stats["fcit_edges_validated"], stats["fcit_sweeps"] = (float(v) for v in st4)
        self._halton_skip += int(st4[0])
        if n.value < 2:
            return None
        path = out[: n.value].astype(np.float64)
        path[0], path[-1, :2], path[-1, 3:] = s, g[:2], g[3:]
        if not cfg.fcit_optimize:
            return path
        return self._shortcut(path, cfg.fcit_reverse_weight)

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
