#!/usr/bin/env python
"""Caller-side latency of the planner surface: small-batch host calls and plan_arm /
plan_whole_body on the synthetic tabletop scene.  One JSON line.  Run on one GPU: python tools/bench_planner.py"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tygrit_b200 import B200PlannerConfig, SE2Pose, WholeBodyConfig, synth  # noqa: E402
from tygrit_b200.planner import create_planner  # noqa: E402


def main():
    planner = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 18))
    planner.update_environment(synth.tabletop_cloud(50_000), SE2Pose(0.0, 0.0, 0.0))
    out = {}
    # host-call latency by batch size (numpy in, numpy out, synchronous)
    for n in (1, 64, 1024, 16384):
        a0, b0, a1, b1 = synth.random_edges(n, seed=3, base_box=(-1, 1, -1, 1), rrt_like=True)
        a = np.concatenate([b0, a0], axis=1).astype(np.float64)
        b = np.concatenate([b1, a1], axis=1).astype(np.float64)
        for _ in range(5):
            planner._edges_ok(a, b)
        t0 = time.perf_counter()
        for _ in range(50):
            planner._edges_ok(a, b)
        out[f"edges_host_call_us_n{n}"] = (time.perf_counter() - t0) / 50 * 1e6
    wb = WholeBodyConfig(arm_joints=np.array([0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]), base_pose=SE2Pose(-0.6, 0.0, 0.0))
    for _ in range(5):
        planner.validate_config(wb)
    t0 = time.perf_counter()
    for _ in range(200):
        planner.validate_config(wb)
    out["validate_config_us"] = (time.perf_counter() - t0) / 200 * 1e6
    # plan_arm between collision-free arm configurations at a base in front of the table
    planner.set_base_params(SE2Pose(-0.1, 0.0, 0.0))
    arm, _ = synth.random_configs(4000, seed=9)
    base = np.tile(np.array([[-0.1, 0.0, 0.0]], np.float32), (len(arm), 1))
    ok = planner.validate_configs(arm, base)
    ok = np.asarray(ok.cpu() if hasattr(ok, "cpu") else ok).astype(bool)
    free = arm[ok].astype(np.float64)
    times, iters, succ = [], [], 0
    for i in range(0, min(len(free) - 1, 40), 2):
        r = planner.plan_arm(free[i], free[i + 1])
        succ += bool(r.success)
        times.append(r.stats.get("planning_time_ms", float("nan")))
        iters.append(r.stats.get("planning_iterations", float("nan")))
    # whole-body planning in the furnished room, all three algorithms (device-resident: no host<->device copies in the loops)
    import torch

    for algo in ("multilayer_rrtc", "fcit_wb", "rrtc_wb"):
        p2 = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 19, whole_body_algorithm=algo, timeout=20.0))
        p2.update_environment(synth.room_cloud(200_000), SE2Pose(0.0, 0.0, 0.0))
        _, bases, _ = synth.free_base_configs(p2.world, 24, seed=5, device=torch.device("cuda", 0), base_box=(-3.2, 3.2, -2.2, 2.2))
        bases = bases.cpu().numpy().astype(np.float64)
        tuck = np.array([0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0])
        tt, ok_n = [], 0
        for i in range(0, 24, 2):
            r = p2.plan_whole_body(WholeBodyConfig(tuck, SE2Pose(*bases[i])), WholeBodyConfig(tuck, SE2Pose(*bases[i + 1])))
            ok_n += bool(r.success)
            tt.append(r.stats.get("total_planning_time_ms", float("nan")))
        out[f"wb_{algo}_success"] = ok_n
        out[f"wb_{algo}_ms_median"] = float(np.nanmedian(tt))
        out[f"wb_{algo}_ms_max"] = float(np.nanmax(tt))
        del p2
    out.update(plan_arm_queries=len(times), plan_arm_success=succ, plan_arm_ms_median=float(np.median(times)),
               plan_arm_ms_max=float(np.max(times)), plan_arm_iterations_median=float(np.median(iters)),
               free_fraction_at_base=float(ok.mean()))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
