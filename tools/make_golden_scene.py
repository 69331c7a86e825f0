#!/usr/bin/env python
"""Golden vectors for the device-resident Scene: run the REFERENCE PointCloudScene
(/root/reference/TyGrit/belief_state/pointcloud_scene.py) on a seeded 5-frame sequence in every
update mode and store its clouds after each update -> tests/golden/scene.npz.

Runs only in the build container (needs /root/reference); the GPU tests read the committed file.
  PYTHONPATH=/root/reference python tools/make_golden_scene.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from TyGrit.belief_state.config import PointCloudSceneConfig  # noqa: E402
from TyGrit.belief_state.pointcloud_scene import PointCloudScene  # noqa: E402
from TyGrit.kinematics.fetch import constants as rc  # noqa: E402
from TyGrit.kinematics.fetch.camera import compute_camera_pose  # noqa: E402
from TyGrit.kinematics.fetch.fk_numpy import forward_kinematics  # noqa: E402
from TyGrit.kinematics.robot_filter import filter_robot_points  # noqa: E402
from TyGrit.types.geometry import SE2Pose  # noqa: E402
from TyGrit.types.robot import RobotState  # noqa: E402
from TyGrit.types.sensor import SensorSnapshot  # noqa: E402
from TyGrit.utils.transforms import se2_to_matrix  # noqa: E402

H, W, F = 120, 160, 5
K = np.array([[154.10223 * W / 640, 0, W / 2], [0, 154.10223 * W / 640, H / 2], [0, 0, 1.0]])


def robot_filter_fn(points, snapshot):
    rs = snapshot.robot_state
    lp = forward_kinematics(np.asarray(list(rs.planning_joints) + list(rs.head_joints), dtype=np.float64))
    Tb = se2_to_matrix(rs.base_pose.x, rs.base_pose.y, rs.base_pose.theta)
    return filter_robot_points(points, {k: Tb @ v for k, v in lp.items()}, rc.FETCH_SPHERES, 0.08)


def make_sequence():
    rng = np.random.default_rng(11)
    u, v = np.meshgrid(np.arange(W), np.arange(H))
    frames, states = [], []
    for f in range(F):
        # a wavy surface 0.5-2.7 m away; the lower half comes close enough to reach the raised arm
        d = 1.6 + 0.9 * np.sin(u / 23.0 + 0.7 * f) * np.cos(v / 17.0 - 0.3 * f) + 0.2 * rng.standard_normal((H, W)) * 0.05
        d[v > 0.55 * H] = 0.35 + 0.5 * ((v[v > 0.55 * H] - 0.55 * H) / (0.45 * H)) + 0.1 * np.sin(u[v > 0.55 * H] / 9.0 + f)
        d = d.astype(np.float32)
        d[rng.random((H, W)) < 0.04] = 0.0
        d[2, 3 + f] = np.nan
        frames.append(d)
        arm = [0.25, 0.3 + 0.1 * f, -0.2, 0.1, 1.2 - 0.1 * f, 0.0, 0.6, 0.0]
        head = [0.1 * f - 0.2, 0.55]
        base = (0.2 + 0.15 * f, -0.1 + 0.05 * f, 0.2 * f - 0.3)
        states.append((arm, head, base))
    return np.stack(frames), states


def main():
    depth, states = make_sequence()
    snaps, cams = [], []
    for f in range(F):
        arm, head, base = states[f]
        rs = RobotState(base_pose=SE2Pose(*base), planning_joints=tuple(arm), head_joints=tuple(head))
        snaps.append(SensorSnapshot(rgb=None, depth=depth[f], intrinsics=K.copy(), robot_state=rs))
        cams.append(compute_camera_pose(rs))
    rng = np.random.default_rng(3)
    static_map = rng.uniform([-1.0, -1.5, 0.35], [2.5, 1.5, 1.8], (4000, 3)).astype(np.float32)
    out = {"depth": depth, "K": K, "cams": np.stack(cams), "static_map": static_map,
           "arm": np.array([s[0] for s in states]), "head": np.array([s[1] for s in states]), "base": np.array([s[2] for s in states])}
    removed_total = 0
    for mode in ("latest", "accumulated", "combine", "ray_casting", "static"):
        for with_static in (False, True):
            for stride in (1, 4):
                cfg = PointCloudSceneConfig(mode=mode, depth_stride=stride)
                sc = PointCloudScene(cfg, static_map if with_static else None, robot_filter_fn=robot_filter_fn)
                sc_nofilter = PointCloudScene(cfg, None)
                for f in range(F):
                    sc.update(snaps[f], cams[f])
                    tag = f"{mode}_s{int(with_static)}_d{stride}_f{f}"
                    # keep the fixture small: M does not depend on the static map; full-resolution runs
                    # are pinned at the last frame only
                    if not with_static and (stride == 4 or f == F - 1):
                        out["M_" + tag] = sc.current_observations()
                    if stride == 4 and f in (0, F - 1):
                        out["E_" + tag] = sc.get_pointcloud()
                    if mode == "latest" and not with_static:
                        sc_nofilter.update(snaps[f], cams[f])
                        removed_total += len(sc_nofilter.current_observations()) - len(sc.current_observations())
    # the oracle's restatement of the two carve predicates, checked against the reference on this sequence
    from oracle import ingest as oin
    from TyGrit.utils.pointcloud import points_in_frustum_mask

    M = out["M_accumulated_s0_d4_f3"]
    ref_mask = points_in_frustum_mask(M, K, cams[4], (0.2, 3.0))
    assert np.array_equal(oin.frustum_mask(M, K, cams[4], (0.2, 3.0)), ref_mask) and 0 < ref_mask.sum() < len(M)
    sc = PointCloudScene(PointCloudSceneConfig(mode="ray_casting", depth_stride=4), None)
    M = out["M_ray_casting_s0_d4_f3"]
    sc._M = M.copy()
    sc._update_ray_casting(np.empty((0, 3), np.float32), depth[4], K, cams[4], (0.2, 3.0))
    keep = oin.raycast_keep(M, depth[4], K, cams[4], 0.03)
    assert np.array_equal(M[keep], sc._M) and 0 < keep.sum() < len(M)
    out["carve_frustum_in"] = ref_mask
    out["carve_ray_keep"] = keep
    assert removed_total > 200, f"robot filter removed only {removed_total} points: the arm is not in view"
    path = os.path.join(ROOT, "tests", "golden", "scene.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; robot filter removed", removed_total, "points over the latest-mode runs")
    for k in sorted(out):
        if k.startswith("M_") and k.endswith(f"f{F - 1}"):
            print(k, out[k].shape)


if __name__ == "__main__":
    main()
