"""Boundary value types of the planner / scene surface.

Structural twins of TyGrit's frozen dataclasses (``TyGrit/types/geometry.py:9-14``,
``types/robot.py:29-42``, ``types/planning.py:28-32``, ``types/results.py:18-25``,
``types/failures.py:16-22``, ``types/sensor.py:14-21``).  TyGrit's own objects can be
passed wherever these appear: only attribute names are relied upon (the reference's
boundary is ``typing.Protocol`` duck typing, ``docs/architecture.md:16``).
"""

from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import Any, Sequence

import numpy as np


@dataclass(frozen=True)
class SE2Pose:
    x: float
    y: float
    theta: float


@dataclass(frozen=True)
class WholeBodyConfig:
    arm_joints: np.ndarray  # (8,) torso + 7 arm
    base_pose: SE2Pose


@dataclass(frozen=True)
class RobotState:
    base_pose: SE2Pose
    planning_joints: tuple[float, ...]  # 8
    head_joints: tuple[float, ...]      # 2


@dataclass(frozen=True)
class SensorSnapshot:
    rgb: Any
    depth: np.ndarray        # (H, W) metres
    intrinsics: np.ndarray   # (3, 3)
    robot_state: RobotState
    segmentation: Any = None


@dataclass(frozen=True)
class Trajectory:
    arm_path: tuple[np.ndarray, ...]
    base_configs: tuple[SE2Pose, ...]


class PlannerFailure(enum.Enum):
    COLLISION_AT_START = "collision_at_start"
    COLLISION_AT_GOAL = "collision_at_goal"
    NO_PATH_FOUND = "no_path_found"
    TIMEOUT = "timeout"


@dataclass(frozen=True)
class PlanResult:
    success: bool
    trajectory: Trajectory | None = None
    failure: PlannerFailure | None = None
    stats: dict[str, float] = field(default_factory=dict)


def tygrit_types() -> dict[str, Any] | None:
    """TyGrit's own value types when the package is importable (a deployment inside TyGrit), else None."""
    try:
        from TyGrit.types.failures import PlannerFailure as _PF
        from TyGrit.types.geometry import SE2Pose as _SE2
        from TyGrit.types.planning import Trajectory as _TR
        from TyGrit.types.results import PlanResult as _PR
        from TyGrit.types.robot import RobotState as _RS, WholeBodyConfig as _WB
        from TyGrit.types.sensor import SensorSnapshot as _SS
    except Exception:
        return None
    return {"SE2Pose": _SE2, "WholeBodyConfig": _WB, "RobotState": _RS, "SensorSnapshot": _SS, "Trajectory": _TR,
            "PlannerFailure": _PF, "PlanResult": _PR}


# Inside TyGrit the planner returns TyGrit's own PlanResult / PlannerFailure / Trajectory, so enum identity and
# isinstance checks of existing callers survive; standalone, the structural twins above are used.
_TWINS = {"SE2Pose": SE2Pose, "WholeBodyConfig": WholeBodyConfig, "RobotState": RobotState, "SensorSnapshot": SensorSnapshot,
          "Trajectory": Trajectory, "PlannerFailure": PlannerFailure, "PlanResult": PlanResult}
_ext = tygrit_types()
USING_TYGRIT_TYPES = _ext is not None
if _ext is not None:
    globals().update(_ext)


def base_xyt(pose: Any) -> tuple[float, float, float]:
    """``(x, y, theta)`` of anything pose-like (SE2Pose, TyGrit SE2Pose, 3-sequence)."""
    if hasattr(pose, "theta"):
        return float(pose.x), float(pose.y), float(pose.theta)
    x, y, th = pose
    return float(x), float(y), float(th)


def make_trajectory(arm_path: Sequence[Sequence[float]], base_configs: Sequence[Sequence[float]]) -> Trajectory:
    """Raw lists -> Trajectory (cf. ``TyGrit/utils/planning.py:35-42``)."""
    return Trajectory(
        arm_path=tuple(np.asarray(c, dtype=np.float64) for c in arm_path),
        base_configs=tuple(SE2Pose(float(b[0]), float(b[1]), float(b[2])) for b in base_configs),
    )
