"""GPU: the callers next to the hot path — batched self-occlusion check against
the reference's golden verdicts, and batched grasp-candidate validation against the serial loop."""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import fk as ofk
from tygrit_b200 import B200PlannerConfig, SE2Pose, WholeBodyConfig, synth
from tygrit_b200.fetch_model import LINK_NAMES
from tygrit_b200.grasp import BatchedGraspValidator, first_valid, validate_grasp_candidates
from tygrit_b200.occlusion import check_self_occlusion, check_self_occlusion_batch
from tygrit_b200.planner import create_planner

pytestmark = pytest.mark.gpu


def test_occlusion_batch_vs_reference_golden(golden):
    g = golden("occlusion.npz")
    got = check_self_occlusion_batch(g["q10"], g["target"], g["radius"], g["K"], tuple(g["shape"])).cpu().numpy()
    assert np.array_equal(got, g["occluded"])                       # 768 states, both verdicts occur
    assert 20 < got.sum() < 700
    n = len(g["occluded_small"])
    small = check_self_occlusion_batch(g["q10"][:n], g["target"][:n], g["radius"][:n], g["K_small"], tuple(g["shape_small"]))
    assert np.array_equal(small.cpu().numpy(), g["occluded_small"])
    # one target for all states
    one = check_self_occlusion_batch(g["q10"][:64], g["target"][3], 0.08, g["K"], tuple(g["shape"])).cpu().numpy()
    each = np.array([check_self_occlusion_batch(g["q10"][i:i + 1], g["target"][3], 0.08, g["K"], tuple(g["shape"])).item() for i in range(64)])
    assert np.array_equal(one, each)


def test_fk_frames_f64_within_1e12_of_reference(golden):
    from tygrit_b200 import _native as N

    g = golden("fk.npz")
    q = torch.from_numpy(np.ascontiguousarray(g["q10"], dtype=np.float64)).cuda()
    out = torch.empty(q.shape[0], 15, 3, 4, dtype=torch.float64, device="cuda")
    N.check(N.lib().tg_fk_frames_f64(q.data_ptr(), q.shape[0], out.data_ptr(), None), "tg_fk_frames_f64")
    np.testing.assert_allclose(out.cpu().numpy(), g["frames"][:, :, :3, :], atol=1e-12)


def test_check_self_occlusion_reference_signature(golden):
    """Dict-of-poses form (checker/occlusion.py:48-66), incl. the missing-head_tilt_link early out."""
    g = golden("occlusion.npz")
    K, shape = g["K"], tuple(int(v) for v in g["shape"])
    idx = np.concatenate([np.nonzero(g["occluded"])[0][:8], np.nonzero(~g["occluded"])[0][:8]])
    for i in idx:
        T = ofk.link_frames_4x4(g["q10"][i:i + 1])[0]
        poses = {name: T[k] for k, name in enumerate(LINK_NAMES)}
        assert check_self_occlusion(poses, g["target"][i], float(g["radius"][i]), K, shape) == bool(g["occluded"][i])
        poses.pop("head_tilt_link")
        assert check_self_occlusion(poses, g["target"][i], float(g["radius"][i]), K, shape) is False


def test_first_valid_groups():
    rng = np.random.default_rng(3)
    v = (rng.random(10_007) < 0.01).astype(np.uint8)
    for group in (1, 7, 200, 10_007, 50_000):
        got = first_valid(torch.from_numpy(v).cuda(), group).cpu().numpy()
        want = []
        for s in range(0, len(v), group):
            nz = np.nonzero(v[s:s + group])[0]
            want.append(nz[0] if len(nz) else -1)
        assert np.array_equal(got, np.array(want, dtype=np.int32))
    assert first_valid(torch.zeros(0, dtype=torch.uint8, device="cuda"), 4).numel() == 0
    assert int(first_valid(torch.from_numpy(v).cuda()).item()) == int(np.nonzero(v)[0][0])


def test_grasp_candidates_batched_equals_serial_loop():
    """The winner of ONE batched launch == the winner of the reference's serial loop over the same IK solutions
    (grasp_sampler.py:109-137), with IK failures (ValueError -> NaN rows) in the mix."""
    planner = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 18))
    planner.update_environment(synth.tabletop_cloud(50_000), SE2Pose(0.0, 0.0, 0.0))
    bp = SE2Pose(0.1, -0.2, 0.3)
    rng = np.random.default_rng(11)
    arm, _ = synth.random_configs(20 * 10, seed=5)
    J = arm.astype(np.float64).reshape(20, 10, 8)
    J[rng.random((20, 10)) < 0.3] = np.nan
    serial = None
    for g in range(20):
        for a in range(10):
            if np.isnan(J[g, a]).any():
                continue
            if planner.validate_config(WholeBodyConfig(arm_joints=J[g, a].astype(np.float32).astype(np.float64), base_pose=bp)):
                serial = (g, a)
                break
        if serial:
            break
    assert validate_grasp_candidates(planner, J, bp) == serial
    assert serial is not None and serial != (0, 0)                  # the cloud rejects the first candidates
    # class form with an injected IK that replays the same solutions
    it = iter(J.reshape(-1, 8))

    def ik(T, seed):
        q = next(it)
        if np.isnan(q).any():
            raise ValueError("no IK solution")
        return q

    out = BatchedGraspValidator(planner, ik, 20, 10).select([np.eye(4)] * 20, bp, seed=np.zeros(8))
    assert np.array_equal(out, J[serial])
    J[:] = np.nan
    assert validate_grasp_candidates(planner, J, bp) is None
