"""CPU: the oracle against the committed golden vectors (made by the reference's own Python,
tools/make_golden.py) and against the reference tests' known-answer values."""

from __future__ import annotations

import numpy as np
import pytest

from oracle import cchecker, collide, fetch_model as om, fk as ofk, ingest as oin
from tygrit_b200 import fetch_model as pm, synth


def test_model_tables_match_reference(golden):
    g = golden("model.npz")
    assert np.array_equal(g["sphere_table"], om.SPHERE_LOCAL)
    assert np.array_equal(g["sphere_table"], pm.SPHERE_XYZR)            # product table too
    assert list(g["sphere_links"]) == list(om.SPHERES_BY_LINK.keys())
    assert [pm.LINK_NAMES[i] for i in pm.SPHERE_LINK] == [om.LINK_NAMES[i] for i in om.SPHERE_LINK]
    assert g["sphere_counts"].sum() == 111
    assert np.array_equal(g["joint_lower"], om.JOINT_LOWER) and np.array_equal(g["joint_upper"], pm.JOINT_UPPER)
    # product kinematic tree == oracle tree
    for (n1, p1, o1, k1, q1), (n2, p2, o2, k2, q2) in zip(pm.LINKS, om.TREE):
        assert n1 == n2 and p1 == p2 and q1 == q2 and np.allclose(o1, o2, atol=0)
        assert {0: "-", 1: "pz", 2: "rx", 3: "ry", 4: "rz"}[k1] == k2


def test_fk_frames_centres_camera_match_reference(golden):
    g = golden("fk.npz")
    np.testing.assert_allclose(ofk.link_frames_4x4(g["q10"]), g["frames"], atol=1e-12)
    np.testing.assert_allclose(ofk.sphere_centres(g["q10"][:, :8], g["base"], g["q10"][:, 8:]), g["centres"], atol=1e-12)
    np.testing.assert_allclose(ofk.camera_pose(g["q10"][:, :8], g["q10"][:, 8:], g["base"]), g["cam2world"], atol=1e-12)


def test_fk_known_answers_sapien():
    """test/test_kinematics.py:208-260 — SAPIEN fixtures 'zeros' and 'head_pan_only', atol 1e-4."""
    T = ofk.link_frames_4x4(np.zeros((1, 10)))[0]
    np.testing.assert_allclose(T[om.LINK_INDEX["gripper_link"], :3, 3], [1.128100037574768, 0, 0.7860098481178284], atol=1e-4)
    np.testing.assert_allclose(T[om.LINK_INDEX["head_tilt_link"], :3, 3], [0.10878008604049683, 0, 1.0384304523468018], atol=1e-4)
    q = np.zeros((1, 10))
    q[0, 8] = 0.5
    T = ofk.link_frames_4x4(q)[0][om.LINK_INDEX["head_tilt_link"]]
    np.testing.assert_allclose(T[:3, 3], [0.09133186936378479, 0.06833262741565704, 1.0384306907653809], atol=1e-4)
    np.testing.assert_allclose(T[0, :2], [0.8775825500488281, -0.4794255495071411], atol=1e-4)
    cam = ofk.camera_pose(np.zeros((1, 8)), [[0.5, 0.0]], np.zeros((1, 3)))[0]
    np.testing.assert_allclose(cam[:3, 3], [0.13959893584251404, 0.09470103681087494, 1.060930609703064], atol=1e-4)


def test_self_filter_matches_reference(golden):
    g = golden("self_filter.npz")
    for i in range(int(g["n_cases"])):
        c32 = oin.filter_centres_f32(g[f"q{i}"][None], g[f"base{i}"][None])[0]
        keep = oin.self_filter_keep(g[f"pts{i}"], c32, float(g[f"radius{i}"]))
        assert np.array_equal(keep, g[f"keep{i}"])
        assert 0 < keep.sum() < len(keep)


def test_self_filter_known_answers():
    """test/test_robot_filter.py:47-105: points at link origins are removed, far points kept."""
    q, b = np.zeros((1, 10)), np.zeros((1, 3))
    c32 = oin.filter_centres_f32(q, b)[0]
    T = ofk.link_frames_4x4(q)[0]
    pts = np.array([T[om.LINK_INDEX["gripper_link"], :3, 3], T[om.LINK_INDEX["elbow_flex_link"], :3, 3], [5, 5, 1]], dtype=np.float32)
    assert list(oin.self_filter_keep(pts, c32, 0.12)) == [False, False, True]
    assert list(oin.self_filter_keep(np.array([[0, 0, 0.2], [5, 5, 1]], np.float32), c32, 0.15)) == [False, True]
    far = oin.filter_centres_f32(q, np.array([[10.0, 10.0, 0.0]]))[0]
    assert oin.self_filter_keep(np.array([[0, 0, 0.2]], np.float32), far, 0.15)[0]


def test_ingest_matches_reference(golden):
    g = golden("ingest.npz")
    for stride in (1, 4):
        pw, keep = oin.depth_to_world(g["depth"], g["K"], g["T_wc"], (0.2, 3.0), stride)
        assert np.array_equal(keep, g[f"keep_s{stride}"])
        np.testing.assert_allclose(pw, g[f"world_s{stride}"], atol=2e-6)
    pc = g["world_s1"]
    assert np.array_equal(oin.crop_xy(pc, g["T_wc"][:3, 3].astype(np.float32), 2.5), g["crop_mask"])
    assert np.array_equal(oin.ground_keep(pc, 0.3), g["ground_mask"])
    assert np.array_equal(oin.voxel_first_index(pc, 0.05), g["voxel_first"])
    assert np.array_equal(oin.merge_add_index(pc[: len(pc) // 2], pc[len(pc) // 2:], 0.03), g["merge_add"])


def test_c_oracle_agrees_with_numpy_oracle():
    cloud = synth.tabletop_cloud(20_000)
    arm, base = synth.random_configs(512, base_box=(-1, 1, -1, 1))
    cl = collide.CloudOracle(cloud).config_clearance(arm, base)
    cc = cchecker.CCloud(cloud)
    np.testing.assert_allclose(cc.clearance(arm, base, cap=10.0), cl, atol=1e-12)
    np.testing.assert_allclose(cchecker.centres_f64(arm, base), ofk.sphere_centres(arm, base), atol=1e-12)
    v32 = cc.check_configs_f32(arm, base).astype(bool)
    band = np.abs(cl) > collide.BAND
    assert np.array_equal(v32[band], (cl > 0)[band])
    assert 0.2 < (cl > 0).mean() < 0.9


def test_edge_sampling_contract():
    a = np.zeros((1, 11))
    b = np.zeros((1, 11))
    b[0, 3] = 0.3                                   # |delta| = 0.3 -> ceil(1.2) = 2 blocks -> 16 samples
    n = collide.edge_n(a, b)
    assert n[0] == 2
    s = collide.edge_samples(a[0], b[0], 2)
    assert s.shape == (16, 11) and np.isclose(s[-1, 3], 0.3) and np.isclose(s[0, 3], 0.3 / 16)
    assert collide.edge_n(a, a)[0] == 1             # zero-length edge still checks the end point (8 samples)
    a0, b0, a1, b1 = synth.random_edges(2000)
    nn = collide.edge_n(collide.edge_vectors(a0, b0), collide.edge_vectors(a1, b1))
    assert 160 < (8 * nn).mean() < 182              # ~171 checks/edge for i.i.d. endpoints


def test_c_oracle_edges_agree():
    cloud = synth.tabletop_cloud(10_000)
    a0, b0, a1, b1 = synth.random_edges(40, base_box=(-1, 1, -1, 1))
    ok, cl, ns = collide.CloudOracle(cloud).edge_valid(a0, b0, a1, b1)
    cc = cchecker.CCloud(cloud)
    cl2, ns2 = cc.edge_clearance(a0, b0, a1, b1, cap=10.0)
    np.testing.assert_allclose(cl2, cl, atol=1e-12)
    assert np.array_equal(ns, ns2)
    v32, total = cc.check_edges_f32(a0, b0, a1, b1)
    assert total == ns.sum()
    band = np.abs(cl) > collide.BAND
    assert np.array_equal(v32.astype(bool)[band], ok[band])


@pytest.mark.parametrize("n", [0, 1])
def test_oracle_degenerate_clouds(n):
    cloud = np.array([[1.0, 0.0, 0.5]], dtype=np.float32)[:n]
    arm, base = synth.random_configs(8)
    cl = collide.CloudOracle(cloud).config_clearance(arm, base)
    assert cl.shape == (8,)
    if n == 0:
        assert np.all(np.isinf(cl))


def test_occlusion_oracle_matches_reference_golden(golden):
    """oracle/occlusion.py vs the unmodified reference's check_self_occlusion (cv2 discs) on seeded states:
    tools/make_golden_occlusion.py; a sample here, the GPU test runs all of them."""
    from oracle import occlusion as oocc

    g = golden("occlusion.npz")
    K, shape = g["K"], tuple(int(v) for v in g["shape"])
    pos = np.nonzero(g["occluded"])[0][:12]
    neg = np.nonzero(~g["occluded"])[0][:12]
    for i in np.concatenate([pos, neg]):
        assert oocc.check_self_occlusion(g["q10"][i], g["target"][i], float(g["radius"][i]), K, shape) == bool(g["occluded"][i])
    Ks, ss = g["K_small"], tuple(int(v) for v in g["shape_small"])
    for i in range(24):
        assert oocc.check_self_occlusion(g["q10"][i], g["target"][i], float(g["radius"][i]), Ks, ss) == bool(g["occluded_small"][i])


def test_disc_rasteriser_known_answer_and_cv2():
    """cv2.circle(mask, c, 5, 1, -1) fills rows of half-width 5,4,4,4,3,0 (81 pixels) — checked against cv2 itself
    wherever cv2 is importable (build container)."""
    from oracle import occlusion as oocc

    assert list(oocc.disc_half_widths(5)) == [5, 4, 4, 4, 3, 0]
    m = np.zeros((20, 20), np.uint8)
    oocc.draw_disc(m, 10, 10, 5)
    assert m.sum() == 81
    cv2 = pytest.importorskip("cv2")
    for r in (1, 2, 3, 7, 31, 64, 200):
        for c in ((40, 30), (2, 58), (79, 0)):
            ref = np.zeros((60, 80), np.uint8)
            cv2.circle(ref, c, r, 1, -1)
            got = np.zeros((60, 80), np.uint8)
            oocc.draw_disc(got, c[0], c[1], r)
            assert np.array_equal(ref, got)


def test_coarse_to_fine_cpu_port_equals_naive_form():
    """The timed CPU baseline (link by link, enclosing sphere first, early exit) gives the verdicts of the naive form
    (all 111 centres, table order) and agrees with the fp64 clearance outside the band."""
    from oracle import cchecker
    from tygrit_b200 import synth

    cloud = synth.room_cloud(60_000)
    cc = cchecker.CCloud(cloud)
    arm, base = synth.random_configs(30_000, seed=11)
    staged = cc.check_configs_f32(arm, base)
    naive = cc.check_configs_f32(arm, base, naive=True)
    assert np.array_equal(staged, naive) and 0.05 < staged.mean() < 0.6
    cl = cc.clearance(arm, base, cap=1e-3)
    d = np.abs(cl) > 1e-5
    assert np.array_equal(staged.astype(bool)[d], (cl > 0)[d])
    one = cc.check_configs_f32(arm[:2000], base[:2000], threads=1)
    assert np.array_equal(one, staged[:2000])
    assert cchecker.set_threads(2) == 2 and cchecker.set_threads(cchecker.host_threads()) == cchecker.host_threads()


def test_self_collision_oracle():
    """Self-collision term: pair list = TyGrit/rl/train.py:78-107 (checked against the reference file when present),
    numpy and C restatements agree, and the reference tests' tuck configuration stays valid."""
    import os

    from oracle import cchecker, collide, fetch_model as fm
    from tygrit_b200 import fetch_model as pm, synth

    assert set(fm.SELF_COLLISION_PAIRS) == set(pm.SELF_COLLISION_PAIRS) and len(fm.SELF_COLLISION_PAIRS) == 26
    ref = "/root/reference/TyGrit/rl/train.py"
    if os.path.exists(ref):
        import ast

        tree = ast.parse(open(ref).read())
        pairs = None
        for node in ast.walk(tree):
            tgt = getattr(node, "target", None) or (node.targets[0] if isinstance(node, ast.Assign) else None)
            if isinstance(node, (ast.Assign, ast.AnnAssign)) and getattr(tgt, "id", "") == "SELF_COLLISION_PAIRS":
                pairs = ast.literal_eval(node.value)
        assert pairs is not None and set(map(tuple, pairs)) == set(fm.SELF_COLLISION_PAIRS)
    arm, _ = synth.random_configs(3000, seed=5)
    a = cchecker.self_clearance(arm)
    b = collide.self_clearance(arm)
    np.testing.assert_allclose(a, b, atol=1e-12)
    assert 0.05 < (a <= 0).mean() < 0.3
    assert collide.self_clearance(np.array([synth.TUCK_ARM]))[0] > 0.005
    cc = cchecker.CCloud(np.empty((0, 3), np.float32))
    v = cc.check_configs_f32(arm, np.zeros((len(arm), 3), np.float32), self_collision=True)
    d = np.abs(a) > 1e-5
    assert np.array_equal(v.astype(bool)[d], (a > 0)[d])
