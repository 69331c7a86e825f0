"""tygrit_b200 — B200-native collision-checking engine behind TyGrit's planner / checker / scene API.

The package holds only what the hot path needs: the CUDA kernels + C ABI (``csrc/``, built into
``lib/libtygrit_b200.so``) and the host-side mirror of the reference interface
(``planner.B200Planner``, ``robot_filter.filter_robot_points``, ``fk.B200BatchFK``,
``scene.B200Scene`` + ``ingest`` for the device-resident cloud ingest and belief merge).
Importing the package does not load CUDA; the first call into the engine does, and fails loudly
if the library is missing (there is no CPU fallback).
"""

from .config import B200PlannerConfig, PlannerConfig  # noqa: F401
from .types import (  # noqa: F401
    PlannerFailure,
    PlanResult,
    RobotState,
    SE2Pose,
    SensorSnapshot,
    Trajectory,
    WholeBodyConfig,
)

__version__ = "0.1.0"


def __getattr__(name):  # lazy: torch / CUDA only when the engine is touched
    if name in ("B200Planner", "create_planner", "create_fetch_planner", "create_collision_check"):
        from . import planner

        return getattr(planner, name)
    if name in ("CollisionWorld",):
        from . import engine

        return getattr(engine, name)
    if name in ("filter_robot_points", "filter_robot_points_batch"):
        from . import robot_filter

        return getattr(robot_filter, name)
    if name in ("B200BatchFK", "batch_forward_kinematics", "create_fk_solver"):
        from . import fk

        return getattr(fk, name)
    if name in ("B200Scene", "PointCloudSceneConfig", "SceneConfig"):
        from . import scene

        return getattr(scene, name)
    if name in ("check_self_occlusion", "check_self_occlusion_batch"):
        from . import occlusion

        return getattr(occlusion, name)
    if name in ("BatchedGraspValidator", "validate_grasp_candidates", "first_valid"):
        from . import grasp

        return getattr(grasp, name)
    raise AttributeError(name)
