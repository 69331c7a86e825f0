#!/usr/bin/env python
"""Generate ``tests/golden/*.npz`` by running the REFERENCE's own Python here.

Run in the build container only (``/root/reference`` does not exist on the GPU
box):  ``python tools/make_golden.py``

It imports the unmodified reference (``TyGrit.kinematics.fetch.fk_numpy``,
``fk_torch``, ``robot_filter``, ``camera``, ``utils.depth``,
``utils.pointcloud``, ``constants``), evaluates it on seeded inputs and stores
inputs + outputs.  It also asserts, on the spot, that the oracle restatement
under ``oracle/`` reproduces every vector — that is the "pin" of the oracle.
"""

from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("TYGRIT_REFERENCE", "/root/reference"))

from TyGrit.kinematics.fetch import constants as rc  # noqa: E402
from TyGrit.kinematics.fetch.camera import compute_camera_pose  # noqa: E402
from TyGrit.kinematics.fetch.fk_numpy import forward_kinematics  # noqa: E402
from TyGrit.kinematics.robot_filter import filter_robot_points  # noqa: E402
from TyGrit.types.geometry import SE2Pose  # noqa: E402
from TyGrit.types.robot import RobotState  # noqa: E402
from TyGrit.utils.depth import depth_to_world_pointcloud  # noqa: E402
from TyGrit.utils.pointcloud import crop_sphere, filter_ground, merge_dedup, voxel_downsample  # noqa: E402
from TyGrit.utils.transforms import se2_to_matrix  # noqa: E402

from oracle import fetch_model as om  # noqa: E402
from oracle import fk as ofk  # noqa: E402
from oracle import ingest as oin  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def ref_world_poses(q10, base):
    lp = forward_kinematics(np.asarray(q10, dtype=np.float64))
    Tb = se2_to_matrix(*base)
    return {k: Tb @ v for k, v in lp.items()}


def ref_sphere_centres(q10, base):
    """World sphere centres the way robot_filter.py:35-44 builds them (fp64)."""
    poses = ref_world_poses(q10, base)
    cs = []
    for link, sph in rc.FETCH_SPHERES.items():
        T = poses[link]
        cs.append(np.asarray(sph, dtype=np.float64)[:, :3] @ T[:3, :3].T + T[:3, 3])
    return np.vstack(cs)


def main() -> None:
    # ---- model table -----------------------------------------------------
    names = list(rc.FETCH_SPHERES.keys())
    table = np.vstack([np.asarray(v, dtype=np.float64) for v in rc.FETCH_SPHERES.values()])
    counts = np.array([len(v) for v in rc.FETCH_SPHERES.values()])
    assert np.array_equal(table, om.SPHERE_LOCAL), "oracle sphere table differs from reference"
    assert names == list(om.SPHERES_BY_LINK.keys())
    np.savez(
        os.path.join(OUT, "model.npz"),
        sphere_table=table, sphere_counts=counts, sphere_links=np.array(names),
        joint_lower=rc.JOINT_LIMITS_LOWER, joint_upper=rc.JOINT_LIMITS_UPPER,
    )
    assert np.array_equal(rc.JOINT_LIMITS_LOWER, om.JOINT_LOWER) and np.array_equal(rc.JOINT_LIMITS_UPPER, om.JOINT_UPPER)

    # ---- FK: link frames, sphere centres, camera pose ----------------------
    rng = np.random.default_rng(42)
    B = 64
    q10 = np.concatenate(
        [rng.uniform(om.JOINT_LOWER, om.JOINT_UPPER, (B, 8)), rng.uniform(om.HEAD_LOWER, om.HEAD_UPPER, (B, 2))], axis=1
    )
    q10[0] = 0.0                                   # the zero pose is a known-answer case
    base = np.stack([rng.uniform(-3, 3, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], axis=1)
    base[0] = 0.0
    frames = np.zeros((B, 15, 4, 4))
    centres = np.zeros((B, 111, 3))
    cams = np.zeros((B, 4, 4))
    link_names = None
    for i in range(B):
        lp = forward_kinematics(q10[i])
        link_names = list(lp.keys())
        for j, n in enumerate(om.LINK_NAMES):
            frames[i, j] = lp[n]
        centres[i] = ref_sphere_centres(q10[i], base[i])
        st = RobotState(base_pose=SE2Pose(*base[i]), planning_joints=tuple(q10[i, :8]), head_joints=tuple(q10[i, 8:]))
        cams[i] = compute_camera_pose(st)
    assert set(link_names) == set(om.LINK_NAMES)
    np.testing.assert_allclose(ofk.link_frames_4x4(q10), frames, atol=1e-12)
    np.testing.assert_allclose(ofk.sphere_centres(q10[:, :8], base, q10[:, 8:]), centres, atol=1e-12)
    np.testing.assert_allclose(ofk.camera_pose(q10[:, :8], q10[:, 8:], base), cams, atol=1e-12)
    # reference known answers (test/test_kinematics.py:214-231)
    np.testing.assert_allclose(frames[0, om.LINK_INDEX["gripper_link"], :3, 3], [1.1281, 0, 0.78601], atol=1e-4)
    np.testing.assert_allclose(frames[0, om.LINK_INDEX["head_tilt_link"], :3, 3], [0.10878, 0, 1.03843], atol=1e-4)
    np.savez(os.path.join(OUT, "fk.npz"), q10=q10, base=base, link_names=np.array(om.LINK_NAMES),
             frames=frames, centres=centres, cam2world=cams)

    # ---- self-filter -------------------------------------------------------
    rng = np.random.default_rng(42)
    cases = {}
    for ci, (q, b, n, rad) in enumerate(
        [(np.zeros(10), (0.0, 0.0, 0.0), 10_000, 0.08),
         (q10[5], tuple(base[5]), 20_000, 0.08),
         (q10[9], tuple(base[9]), 20_000, 0.12)]
    ):
        poses = ref_world_poses(q, b)
        far = rng.uniform(-5, 5, (n // 2, 3))
        cen = ref_sphere_centres(q, b)
        near = cen[rng.integers(0, 111, n - n // 2)] + rng.normal(0, 0.07, (n - n // 2, 3))
        pts = np.vstack([far, near]).astype(np.float32)
        kept = filter_robot_points(pts, poses, rc.FETCH_SPHERES, sphere_radius=rad)
        # recover the keep mask from the kept rows (order preserved)
        c32 = oin.filter_centres_f32(q[None], np.asarray(b)[None])[0]
        keep = oin.self_filter_keep(pts, c32, rad)
        assert np.array_equal(pts[keep], kept), f"self-filter oracle mismatch case {ci}"
        cases[f"q{ci}"], cases[f"base{ci}"], cases[f"pts{ci}"] = np.asarray(q), np.asarray(b), pts
        cases[f"keep{ci}"], cases[f"radius{ci}"] = keep, np.float64(rad)
    np.savez_compressed(os.path.join(OUT, "self_filter.npz"), n_cases=3, **cases)

    # ---- depth ingest ------------------------------------------------------
    rng = np.random.default_rng(7)
    H, W = 96, 128
    K = np.array([[154.10223 * W / 640, 0, W / 2], [0, 154.10223 * W / 640, H / 2], [0, 0, 1.0]])
    depth = rng.uniform(0.1, 3.5, (H, W)).astype(np.float32)
    depth[rng.random((H, W)) < 0.05] = 0.0
    depth[3, 5] = np.nan
    T = cams[7]
    out = {}
    for stride in (1, 4):
        ref = depth_to_world_pointcloud(depth, K, T, z_range=(0.2, 3.0), stride=stride)
        mine, keep = oin.depth_to_world(depth, K, T, (0.2, 3.0), stride)
        assert mine.shape == ref.shape
        np.testing.assert_allclose(mine, ref, atol=2e-6)
        out[f"world_s{stride}"] = ref
        out[f"keep_s{stride}"] = keep
    pc = out["world_s1"]
    cm = oin.crop_xy(pc, T[:3, 3].astype(np.float32), 2.5)
    assert np.array_equal(pc[cm], crop_sphere(pc, T[:3, 3].astype(np.float32), 2.5))
    gm = oin.ground_keep(pc, 0.3)
    assert np.array_equal(pc[gm], filter_ground(pc, 0.3))
    vi = oin.voxel_first_index(pc, 0.05)
    assert np.array_equal(pc[vi], voxel_downsample(pc, 0.05))
    half = len(pc) // 2
    ai = oin.merge_add_index(pc[:half], pc[half:], 0.03)
    assert np.array_equal(np.vstack((pc[:half], pc[half:][ai])), merge_dedup(pc[:half], pc[half:], 0.03))
    np.savez_compressed(os.path.join(OUT, "ingest.npz"), depth=depth, K=K, T_wc=T, crop_mask=cm, ground_mask=gm,
                        voxel_first=vi, merge_add=ai, **out)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
