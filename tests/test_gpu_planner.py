"""GPU: the MotionPlanner / CollisionCheckFn surface, mirroring the reference's test/test_planner.py
(which is skipped upstream wherever VAMP is absent) plus verdict checks against the oracle."""

from __future__ import annotations

import numpy as np
import pytest

from oracle import collide
from tygrit_b200 import B200PlannerConfig, PlannerFailure, SE2Pose, WholeBodyConfig, synth
from tygrit_b200.planner import create_collision_check, create_planner

pytestmark = pytest.mark.gpu

_TUCK = [0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]
_READY = [0.05, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0]


@pytest.fixture(scope="module")
def planner():
    return create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 18))


def test_validate_config_tuck_empty_env(planner):
    planner.clear_pointclouds()
    wb = WholeBodyConfig(arm_joints=np.array(_TUCK), base_pose=SE2Pose(0.0, 0.0, 0.0))
    assert planner.validate_config(wb) is True                      # test/test_planner.py:72-78
    assert create_collision_check(planner)(wb) is True


def test_plan_arm_empty_env(planner):
    planner.clear_pointclouds()
    res = planner.plan_arm(np.array(_TUCK), np.array(_READY))
    assert res.success and res.trajectory is not None and len(res.trajectory.arm_path) >= 2
    for cfg in res.trajectory.arm_path:
        assert cfg.shape == (8,)
    assert res.trajectory.base_configs[0] == SE2Pose(0.0, 0.0, 0.0)
    assert "planning_time_ms" in res.stats


def test_add_clear_pointcloud_and_spheres(planner):
    planner.add_pointcloud(np.array([[1.0, 0.0, 0.5], [1.1, 0.1, 0.5]], dtype=np.float64))
    planner.clear_pointclouds()
    wb = WholeBodyConfig(np.array(_READY), SE2Pose(0.0, 0.0, 0.0))
    assert planner.validate_config(wb)
    planner.add_sphere([0.9, 0.0, 0.75], 0.1)                        # on the stretched arm
    assert not planner.validate_config(wb)
    planner.clear_spheres()
    assert planner.validate_config(wb)


def test_attach_detach(planner):
    spheres = [{"position": [0.0, 0.0, 0.0], "radius": 0.02}]
    assert planner.attach_to_eef(spheres) is True
    assert planner.attach_to_eef(spheres) is False
    assert planner.detach_from_eef() is True
    assert planner.detach_from_eef() is False


def test_update_environment_and_verdicts_match_oracle(planner):
    cloud = synth.tabletop_cloud(50_000)
    planner.update_environment(cloud, SE2Pose(0.0, 0.0, 0.0))
    arm, base = synth.random_configs(64, seed=5, base_box=(-1, 1, -1, 1))
    cl = collide.CloudOracle(cloud).config_clearance(arm, base)
    for i in range(64):
        if abs(cl[i]) > 1e-5:
            wb = WholeBodyConfig(arm[i].astype(np.float64), SE2Pose(*map(float, base[i])))
            assert planner.validate_config(wb) == bool(cl[i] > 0)
    ok = planner.validate_configs(arm, base)
    assert np.array_equal(ok[np.abs(cl) > 1e-5], (cl > 0)[np.abs(cl) > 1e-5])
    planner.update_environment(np.empty((0, 3), np.float32), SE2Pose(0, 0, 0))   # empty cloud just clears
    assert planner.validate_configs(arm, base).all()


def test_plan_interpolation(planner):
    res = planner.plan_interpolation(np.array(_TUCK), np.array(_READY), SE2Pose(1.0, 2.0, 0.5))
    assert res.success and len(res.trajectory.arm_path) == 16
    np.testing.assert_allclose(res.trajectory.arm_path[0], _TUCK)
    np.testing.assert_allclose(res.trajectory.arm_path[-1], _READY)
    assert all(b == SE2Pose(1.0, 2.0, 0.5) for b in res.trajectory.base_configs)


def test_plan_whole_body_and_failures(planner):
    planner.clear_pointclouds()
    start = WholeBodyConfig(np.array(_TUCK), SE2Pose(0.0, 0.0, 0.0))
    goal = WholeBodyConfig(np.array(_TUCK), SE2Pose(0.5, 0.0, 0.0))
    res = planner.plan_whole_body(start, goal)
    assert res.success and len(res.trajectory.base_configs) == len(res.trajectory.arm_path) >= 2
    # wall of points in front of the goal base pose: goal in collision
    wall = np.stack(np.meshgrid(np.array([0.5]), np.linspace(-0.5, 0.5, 40), np.linspace(0.3, 1.2, 40)), -1).reshape(-1, 3)
    planner.add_pointcloud(wall.astype(np.float32))
    res = planner.plan_whole_body(start, goal)
    assert not res.success and res.failure == PlannerFailure.COLLISION_AT_GOAL
    res = planner.plan_whole_body(goal, start)
    assert not res.success and res.failure == PlannerFailure.COLLISION_AT_START
    planner.clear_pointclouds()


def test_plan_arm_around_obstacle_is_collision_free(planner):
    cloud = synth.tabletop_cloud(20_000)
    planner.update_environment(cloud, SE2Pose(0.0, 0.0, 0.0))
    orc = collide.CloudOracle(cloud)
    start = np.array(_TUCK)
    goal = np.array([0.3, 0.6, -0.4, 0.0, 0.9, 0.0, 0.6, 0.0])
    base0 = np.zeros((1, 3))
    if not (orc.config_clearance(start[None], base0)[0] > 0 and orc.config_clearance(goal[None], base0)[0] > 0):
        pytest.skip("fixture configs collide in this cloud")
    res = planner.plan_arm(start, goal)
    assert res.success, res.failure
    path = np.asarray(res.trajectory.arm_path)
    np.testing.assert_allclose(path[0], start, atol=1e-6)
    np.testing.assert_allclose(path[-1], goal, atol=1e-6)
    cl = orc.config_clearance(path, np.zeros((len(path), 3)))
    assert (cl > -1e-5).all()                                        # every waypoint is free per the oracle
    assert planner.check_trajectory_collisions(path.tolist(), [[0, 0, 0]] * len(path)) is False
    bad = path.copy()
    bad[len(bad) // 2] = [0.0, 0.0, 0.9, 0.0, 0.0, 0.0, 0.0, 0.0]     # arm pushed down into the table
    if orc.config_clearance(bad[len(bad) // 2][None], base0)[0] < -1e-3:
        assert planner.check_trajectory_collisions(bad.tolist(), [[0, 0, 0]] * len(bad)) is True
