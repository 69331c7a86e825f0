#!/usr/bin/env python
"""CPU baselines of the reference's OWN in-tree code, imported unmodified from /root/reference.

Runs in the BUILD CONTAINER only (the reference tree does not travel to the GPU box): writes
profiles/r02_cpu_reference_baselines.json.  Each figure says how many host threads it used.

  filter_robot_points (kinematics/robot_filter.py:13-64)     8 head-camera frames, full 640x480 and stride 4
  batch_forward_kinematics (kinematics/fetch/fk_torch.py)     4096 configs, torch CPU threads
  forward_kinematics (kinematics/fetch/fk_numpy.py)           one config at a time
  depth_to_world_pointcloud + voxel_downsample                one frame, stride 1 and 4
"""
import json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
import torch  # noqa: E402
from TyGrit.kinematics.fetch.constants import FETCH_SPHERES  # noqa: E402
from TyGrit.kinematics.fetch.fk_numpy import forward_kinematics  # noqa: E402
from TyGrit.kinematics.fetch.fk_torch import batch_forward_kinematics  # noqa: E402
from TyGrit.kinematics.robot_filter import filter_robot_points  # noqa: E402
from TyGrit.utils.depth import depth_to_world_pointcloud  # noqa: E402
from TyGrit.utils.pointcloud import voxel_downsample  # noqa: E402
from oracle import fk as ofk  # noqa: E402
from tygrit_b200 import synth  # noqa: E402

rng = np.random.default_rng(42)
states = synth.random_states(8)
K = synth.HEAD_CAMERA_K
out = {"host": {"cpu_count": os.cpu_count(), "torch_threads": torch.get_num_threads()}, "where": "build container (no GPU)"}
# synthetic depth: a wall 2 m ahead with noise (the filter's cost does not depend on what the pixels show)
depth = (2.0 + 0.3 * rng.random((480, 640))).astype(np.float32)
t_full, t_s4, t_d2w1, t_d2w4 = [], [], [], []
for s in states:
    T = ofk.camera_pose(s[None, 3:11], s[None, 11:13], s[None, :3])[0]
    poses = forward_kinematics(s[3:13].astype(np.float64))
    c, th = np.cos(s[2]), np.sin(s[2])
    Tb = np.array([[c, -th, 0, s[0]], [th, c, 0, s[1]], [0, 0, 1, 0], [0, 0, 0, 1]])
    poses = {k: Tb @ v for k, v in poses.items()}
    for stride, acc_d, acc_f in ((1, t_d2w1, t_full), (4, t_d2w4, t_s4)):
        t0 = time.perf_counter(); pts = depth_to_world_pointcloud(depth, K, T, z_range=(0.2, 3.0), stride=stride); acc_d.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); filter_robot_points(pts, poses, FETCH_SPHERES); acc_f.append(time.perf_counter() - t0)
out["filter_robot_points"] = {"full_frame_s": float(np.median(t_full)), "stride4_frame_s": float(np.median(t_s4)), "frames": 8, "threads": 1,
                              "extrapolated_1024_full_frames_s": float(np.median(t_full)) * 1024,
                              "cite": "TyGrit/kinematics/robot_filter.py:13-64"}
out["depth_to_world_pointcloud"] = {"full_frame_s": float(np.median(t_d2w1)), "stride4_frame_s": float(np.median(t_d2w4)),
                                    "cite": "TyGrit/utils/depth.py:52-84"}
pts = depth_to_world_pointcloud(depth, K, np.eye(4), z_range=(0.2, 3.0), stride=1)
t0 = time.perf_counter(); voxel_downsample(pts, 0.05); out["voxel_downsample"] = {"s_per_307k_points": time.perf_counter() - t0,
                                                                                 "cite": "TyGrit/utils/pointcloud.py:13-31"}
q = torch.from_numpy(synth.random_states(4096)[:, 3:13].copy())
batch_forward_kinematics(q)
t0 = time.perf_counter()
for _ in range(5):
    batch_forward_kinematics(q)
dt = (time.perf_counter() - t0) / 5
out["batch_forward_kinematics"] = {"configs_per_s": 4096 / dt, "batch": 4096, "threads": torch.get_num_threads(),
                                   "cite": "TyGrit/kinematics/fetch/fk_torch.py:119-283"}
t0 = time.perf_counter()
for i in range(200):
    forward_kinematics(states[i % 8, 3:13].astype(np.float64))
out["forward_kinematics"] = {"configs_per_s": 200 / (time.perf_counter() - t0), "threads": 1, "cite": "TyGrit/kinematics/fetch/fk_numpy.py:32-157"}
path = os.path.join(ROOT, "profiles", "r02_cpu_reference_baselines.json")
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out))
