#!/usr/bin/env python
"""Verdict + exact clearance on the bench cloud (500k-point room, 100 voxels/m): time per batch and the largest difference
from the fp64 oracle on a sample.  TYGRIT_B200_LIB selects a variant build.

  python tools/clearance_time.py [--configs N] [--sample M] [--workload uniform|free]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cchecker  # noqa: E402  (checker only)
from tygrit_b200 import _native as N, synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--configs", type=int, default=1 << 22)
p.add_argument("--sample", type=int, default=1 << 17)
p.add_argument("--vpm", type=float, default=100.0)
p.add_argument("--workload", default="uniform", choices=["uniform", "free"])
a = p.parse_args()

cloud = synth.room_cloud(500_000)
w = CollisionWorld(device=0, max_points=1 << 20, voxels_per_metre=a.vpm, max_voxels=192 << 20)
w.build(torch.from_numpy(cloud).cuda())
dev = torch.device("cuda", 0)
arm, base = synth.device_configs(a.configs, seed=42, device=dev)
if a.workload == "free":
    arm, base, _ = synth.free_base_configs(w, a.configs, seed=42, device=dev)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


v_ms = timed(lambda: w.check_configs(arm, base))
c_ms = timed(lambda: w.check_configs(arm, base, clearance=True))
valid, clr = w.check_configs(arm, base, clearance=True)
n = min(a.sample, a.configs)
ref = cchecker.CCloud(cloud).clearance(arm[:n].cpu().numpy(), base[:n].cpu().numpy(), cap=0.25)
got = clr[:n].cpu().numpy()
err = np.abs(got - ref)
print(json.dumps({
    "lib": os.path.basename(N.LIB_PATH), "workload": a.workload, "configs": a.configs, "verdict_ms": v_ms, "verdict_plus_clearance_ms": c_ms,
    "configs_per_s_verdict": a.configs / (v_ms * 1e-3), "configs_per_s_with_clearance": a.configs / (c_ms * 1e-3),
    "ratio": v_ms / c_ms, "sample": n, "max_abs_err_m": float(err.max()), "n_err_gt_3e-6": int((err > 3e-6).sum()),
    "free_fraction": float(valid.float().mean().item()), "at_cap_fraction": float((got >= 0.25).mean()),
}))
