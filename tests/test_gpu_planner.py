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
    assert "planning_time_ms" in res.stats and "planning_iterations" in res.stats   # the reference's keys (algorithms.py:43-44)


def test_add_clear_pointcloud_and_spheres(planner):
    planner.add_pointcloud(np.array([[1.0, 0.0, 0.5], [1.1, 0.1, 0.5]], dtype=np.float64))
    planner.clear_pointclouds()
    wb = WholeBodyConfig(np.array(_READY), SE2Pose(0.0, 0.0, 0.0))
    assert planner.validate_config(wb)
    planner.add_sphere([0.9, 0.0, 0.75], 0.1)                        # on the stretched arm
    assert not planner.validate_config(wb)
    planner.clear_spheres()
    assert planner.validate_config(wb)


def test_pointclouds_accumulate_until_cleared(planner):
    """vamp_preview.Environment.add_pointcloud ADDS geometry until clear_pointclouds (vamp_preview.py:205-240)."""
    planner.clear_pointclouds()
    wb = WholeBodyConfig(np.array(_READY), SE2Pose(0.0, 0.0, 0.0))
    near_arm = np.array([[0.9, 0.0, 0.78]], dtype=np.float64)          # on the stretched arm
    far = np.array([[3.0, 3.0, 1.0], [3.1, 3.0, 1.0]], dtype=np.float64)
    planner.add_pointcloud(near_arm)
    assert not planner.validate_config(wb)
    planner.add_pointcloud(far)                                        # a second cloud must not drop the first
    assert not planner.validate_config(wb)
    planner.clear_pointclouds()
    planner.add_pointcloud(far)
    assert planner.validate_config(wb)
    planner.clear_pointclouds()


def test_sphere_capacity_checked_before_mutating(planner):
    planner.clear_spheres()
    for i in range(64):
        planner.add_sphere([5.0 + i, 5.0, 1.0], 0.05)
    with pytest.raises(RuntimeError):
        planner.add_sphere([0.9, 0.0, 0.75], 0.1)                      # the 65th: rejected, host list and device in step
    assert len(planner._obstacles) == 64
    assert planner.validate_config(WholeBodyConfig(np.array(_READY), SE2Pose(0.0, 0.0, 0.0)))
    planner.clear_spheres()


def test_unknown_algorithm_names_raise():
    for kw in ({"algorithm": "prm"}, {"whole_body_algorithm": "nope"}, {"sampler": "xorshift"}):
        with pytest.raises(ValueError):
            create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 12, **kw))


def test_self_collision_is_part_of_the_planner_verdict(planner):
    """B200Planner enables the self-collision term (VAMP's validate_whole_body_config tests the robot against itself)."""
    from oracle import cchecker

    planner.clear_pointclouds()
    arm, base = synth.random_configs(4000, seed=77)
    cl = cchecker.self_clearance(arm)
    ok = np.asarray(planner.validate_configs(arm, base))
    d = np.abs(cl) > 1e-5
    assert np.array_equal(ok[d], (cl > 0)[d]) and 0.05 < (~ok).mean() < 0.3
    i = int(np.flatnonzero(cl < -1e-3)[0])
    assert planner.validate_config(WholeBodyConfig(arm[i].astype(np.float64), SE2Pose(0.0, 0.0, 0.0))) is False


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
    from oracle import cchecker

    cl = np.minimum(collide.CloudOracle(cloud).config_clearance(arm, base), cchecker.self_clearance(arm))   # the planner's verdict
    for i in range(64):
        if abs(cl[i]) > 1e-5:
            wb = WholeBodyConfig(arm[i].astype(np.float64), SE2Pose(*map(float, base[i])))
            assert planner.validate_config(wb) == bool(cl[i] > 0)
    ok = planner.validate_configs(arm, base)
    assert np.array_equal(ok[np.abs(cl) > 1e-5], (cl > 0)[np.abs(cl) > 1e-5])
    planner.update_environment(np.empty((0, 3), np.float32), SE2Pose(0, 0, 0))   # empty cloud just clears
    only_self = cchecker.self_clearance(arm)                                       # what is left: the self-collision term
    got = planner.validate_configs(arm, base)
    assert np.array_equal(got[np.abs(only_self) > 1e-5], (only_self > 0)[np.abs(only_self) > 1e-5]) and got.mean() > 0.7


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


def _room_planner(algo):
    p = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 19, whole_body_algorithm=algo, timeout=20.0))
    cloud = synth.room_cloud(150_000)
    p.update_environment(cloud, SE2Pose(0.0, 0.0, 0.0))
    return p, cloud


def _free_base_poses(p, n, seed):
    import torch

    arm, base, _ = synth.free_base_configs(p.world, n, seed=seed, device=torch.device("cuda", 0), base_box=(-3.2, 3.2, -2.2, 2.2))
    return base.cpu().numpy().astype(np.float64)


@pytest.mark.parametrize("algo", ["multilayer_rrtc", "fcit_wb", "rrtc_wb"])
def test_whole_body_algorithms_in_the_room(algo):
    """Every whole_body_algorithm plans between free poses of the furnished room; the returned trajectory is collision-free
    per the ORACLE (cloud + self-collision) at every waypoint, starts / ends at the request and turns along the short arc."""
    from oracle import cchecker

    p, cloud = _room_planner(algo)
    poses = _free_base_poses(p, 16, seed=3)
    cc = cchecker.CCloud(cloud)
    solved = tried = 0
    raised = [0.25, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]                # tuck with the torso up: the arm layer has work to do
    for i in range(0, 16, 2):                                                # some pairs lie in pockets the furniture closes off
        s = WholeBodyConfig(np.array(_TUCK), SE2Pose(*poses[i]))
        g = WholeBodyConfig(np.array(_TUCK if i % 4 else raised), SE2Pose(*poses[i + 1]))
        if not (p.validate_config(s) and p.validate_config(g)):
            continue
        tried += 1
        res = p.plan_whole_body(s, g)
        if not res.success:
            assert res.failure == PlannerFailure.NO_PATH_FOUND
            continue
        solved += 1
        arm = np.asarray(res.trajectory.arm_path)
        base = np.array([[b.x, b.y, b.theta] for b in res.trajectory.base_configs])
        np.testing.assert_allclose(arm[0], s.arm_joints, atol=1e-6)
        np.testing.assert_allclose(arm[-1], g.arm_joints, atol=1e-6)
        np.testing.assert_allclose(base[0], poses[i], atol=1e-6)
        np.testing.assert_allclose(base[-1, :2], poses[i + 1, :2], atol=1e-6)
        assert abs(np.angle(np.exp(1j * (base[-1, 2] - poses[i + 1, 2])))) < 1e-5         # same heading modulo 2 pi
        assert np.abs(np.diff(base[:, 2])).max() < 0.5                                   # no 2 pi jumps between waypoints
        cl = np.minimum(cc.clearance(arm.astype(np.float32), base.astype(np.float32), cap=1e-3), cchecker.self_clearance(arm.astype(np.float32)))
        assert (cl > -1e-5).all(), f"{algo}: waypoint in collision per the oracle (min clearance {cl.min()})"
        assert p.check_trajectory_collisions(arm.tolist(), base.tolist()) is False
        assert "total_planning_time_ms" in res.stats
    assert solved >= 2, f"{algo}: only {solved} of {tried} queries solved"   # 5 of the 8 pairs sit in pockets no planner leaves


def test_multilayer_base_layer_prefers_driving_forward():
    """reverse_penalty = 50 (algorithms.py:93-96): to reach a pose 1.5 m BEHIND it in free space the robot turns round
    rather than reversing all the way."""
    p = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 12))
    p.clear_pointclouds()
    s = WholeBodyConfig(np.array(_TUCK), SE2Pose(0.0, 0.0, 0.0))
    g = WholeBodyConfig(np.array(_TUCK), SE2Pose(-1.5, 0.0, 0.0))
    res = p.plan_whole_body(s, g)
    assert res.success and "base_planning_time_ms" in res.stats
    base = np.array([[b.x, b.y, b.theta] for b in res.trajectory.base_configs])
    d = np.diff(base[:, :2], axis=0)
    heading = np.stack([np.cos(base[:-1, 2]), np.sin(base[:-1, 2])], 1)
    moving = np.linalg.norm(d, axis=1) > 1e-6
    backwards = ((d * heading).sum(1) < -1e-6) & moving
    assert np.linalg.norm(d[backwards], axis=1).sum() < 0.5 * np.linalg.norm(d[moving], axis=1).sum()


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
