#!/usr/bin/env python
"""Secondary benchmark lines for the BASELINE.json configurations that are not bench.py's headline.

  C1  4096 random configs vs 50k-point tabletop cloud          (latency of one call, device + host API)
  C2  100k random arm edges vs 200k-point room cloud            (edges/s, config checks/s)
  C3  self-filter, 640x480 depth x 1024 robot states            (pixels/s, GB/s vs HBM peak, 5 B/pixel)
  C5  100 steps x (cloud rebuild + 1M edges)                    (steps/s, build us, edges/s)

One JSON line per configuration on stdout.  Inputs are synthetic (tygrit_b200/synth.py) and resident
on the device unless the line says "host".  Run on one GPU: python tools/bench_configs.py [--only C2]
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tygrit_b200 import _native as N  # noqa: E402
from tygrit_b200 import engine, synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def cpu_edges(cloud, edges_np, sample=2000):
    """cpu_baseline for the edge lines: the oracle's C port (coarse-to-fine fp32 checker) on a bounded sample of the same
    edges, all host threads and one thread."""
    import time as _t

    from oracle import cchecker

    cc = cchecker.CCloud(cloud)
    e = [x[:sample] for x in edges_np]
    th = cchecker.host_threads()
    cc.check_edges_f32(*[x[:200] for x in e], threads=th)
    t0 = _t.perf_counter(); _, n = cc.check_edges_f32(*e, threads=th); dt = _t.perf_counter() - t0
    m = max(200, sample // 8)
    t1 = _t.perf_counter(); _, n1 = cc.check_edges_f32(*[x[:m] for x in e], threads=1); dt1 = _t.perf_counter() - t1
    cchecker.set_threads(th)
    return {"value": sample / dt, "unit": "edges/s", "config_checks_per_s": n / dt, "cores": th, "kind": "port",
            "sample": f"{sample} of the same edges, oracle/cpu_checker.c fp32 coarse-to-fine checker (early exit per edge), OpenMP {th} threads",
            "one_core": {"value": m / dt1, "unit": "edges/s", "config_checks_per_s": n1 / dt1, "cores": 1, "sample": f"{m} edges"}}


def c1(args):
    cloud = synth.tabletop_cloud(50_000)
    w = CollisionWorld(max_points=1 << 20, voxels_per_metre=args.vpm)
    w.build(cuda(cloud))
    arm, base = synth.random_configs(4096, base_box=(-1, 1, -1, 1))
    a, b = cuda(arm), cuda(base)
    v = torch.empty(4096, dtype=torch.uint8, device="cuda")
    ms = timed(lambda: w.check_configs_into(a, b, v), 200)
    t0 = time.perf_counter()
    for _ in range(50):
        hv = w.check_configs_host(arm, base)
    host_ms = (time.perf_counter() - t0) / 50 * 1e3
    return {"config": "C1: 4096 configs vs 50k-point tabletop cloud", "device_call_us": ms * 1e3, "configs_per_s": 4096 / (ms * 1e-3),
            "host_call_us": host_ms * 1e3, "host_configs_per_s": 4096 / (host_ms * 1e-3), "free_fraction": float(hv.mean())}


def c2(args):
    cloud = synth.room_cloud(200_000)
    w = CollisionWorld(max_points=1 << 20, voxels_per_metre=args.vpm)
    w.build(cuda(cloud))
    out = {}
    for name, rrt in (("uniform", False), ("rrt_like", True)):
        e_np = synth.random_edges(100_000, rrt_like=rrt)
        e = [cuda(x) for x in e_np]
        valid, n = w.check_edges(*e, count=True)
        checks = int(n.item())
        ms = timed(lambda: w.check_edges(*e), 100, warm=100)
        out[name] = {"ms": ms, "edges_per_s": 1e5 / (ms * 1e-3), "config_checks": checks,
                     "config_checks_per_s": checks / (ms * 1e-3), "valid_fraction": float(valid.float().mean().item()),
                     "valid_edges_per_s": float(valid.float().sum().item()) / (ms * 1e-3)}
        if not args.no_cpu:
            out[name]["cpu_baseline"] = cpu_edges(cloud, e_np, 4000 if rrt else 2000)
    return {"config": "C2: 100k random Fetch edges (rake 8, resolution 32) vs 200k-point room cloud", **out}


def c3(args):
    S, H, W = args.states, 480, 640
    states = cuda(synth.random_states(S))
    boxes = synth.room_boxes()
    bx = cuda(np.array([[*lo, *hi] for lo, hi in boxes], np.float32))
    depth = torch.empty(S, H, W, device="cuda")
    K = np.ascontiguousarray(synth.HEAD_CAMERA_K)
    N.check(N.lib().tg_synth_depth(states.data_ptr(), S, bx.data_ptr(), len(boxes), H, W, K.ctypes.data, 0.0, 1,
                                   depth.data_ptr(), None), "tg_synth_depth")
    mask = torch.empty(S, H, W, dtype=torch.uint8, device="cuda")
    # ~100 ms of warm-up so that the clocks have ramped, then the mean of 100 back-to-back calls
    ms = timed(lambda: engine.self_filter_depth(depth, K, states, 0.08, (0.2, 3.0), 1, mask=mask), 100, warm=150)
    px = S * H * W
    gbs = px * 5 / (ms * 1e-3) / 1e9
    mask4 = torch.empty(S, (H + 3) // 4, (W + 3) // 4, dtype=torch.uint8, device="cuda")
    ms4 = timed(lambda: engine.self_filter_depth(depth, K, states, 0.08, (0.2, 3.0), 4, mask=mask4), 100, warm=100)
    valid = depth > 0
    cpu = None
    if not args.no_cpu:
        import time as _t

        from oracle import fk as ofk, ingest as oin

        st_np, d_np = states[:4].cpu().numpy(), depth[:4].cpu().numpy()
        t0 = _t.perf_counter()
        for s in range(4):
            T = ofk.camera_pose(st_np[s:s + 1, 3:11], st_np[s:s + 1, 11:13], st_np[s:s + 1, :3])[0]
            pw, keep = oin.depth_to_world(d_np[s], K, T, (0.2, 3.0), 1)
            c32 = oin.filter_centres_f32(st_np[s:s + 1, 3:13], st_np[s:s + 1, :3])[0]
            oin.self_filter_keep(pw, c32, 0.08)
        dt = (_t.perf_counter() - t0) / 4
        cpu = {"value": 1.0 / dt, "unit": "frames/s", "pixels_per_s": H * W / dt, "cores": 1, "kind": "port",
               "sample": "4 of the states, full 640x480 frames: oracle/ingest.py (numpy restatement of utils/depth.py + "
                         "kinematics/robot_filter.py, single thread like the reference); the reference's own functions timed in the "
                         "build container: profiles/r02_cpu_reference_baselines.json"}
    return {"cpu_baseline": cpu, "config": f"C3: robot self-filter, {W}x{H} depth x {S} states", "ms": ms, "pixels_per_s": px / (ms * 1e-3),
            "frames_per_s": S / (ms * 1e-3), "algorithmic_GBps": gbs, "hbm_peak_GBps": hbm_peak(), "frac_of_hbm": gbs / hbm_peak(),
            "stride4_ms": ms4, "kept_fraction": float(mask.float().mean().item()), "valid_depth_fraction": float(valid.float().mean().item()),
            "robot_pixels_fraction": float((valid & (mask == 0) & (depth >= 0.2) & (depth <= 3.0)).float().mean().item())}


def c5(args):
    """rebuild + edge validation per step: CUDA-graph replays timed from C (tg_c5_run), fresh device-generated edges
    every step; the Python-driven loop of round 1 (same edges every step) is kept beside it."""
    steps, E = args.steps, args.edges
    w = CollisionWorld(max_points=1 << 20, voxels_per_metre=args.vpm)
    w.set_bounds((-4.2, -3.2, 0.2), (4.2, 3.2, 2.8))          # fixed workspace: the build never synchronises
    clouds_np = [synth.room_cloud(200_000 + 100_000 * (i % 4), perturb_step=i) for i in range(8)]
    clouds = [cuda(c) for c in clouds_np]
    runs = [w.run_replanning_steps(clouds, E, steps=steps, warmup=5, seed=42 + r) for r in range(3)]
    g = sorted(runs, key=lambda r: r["ms_per_step"])[1]      # median of three runs
    spread = (max(r["ms_per_step"] for r in runs) - min(r["ms_per_step"] for r in runs)) / g["ms_per_step"]
    out = {"config": f"C5: {steps} steps x (rebuild 200k-500k-point cloud + {E} fresh edges), one CUDA graph per step, timed from C",
           "ms_per_step": g["ms_per_step"], "steps_per_s": 1e3 / g["ms_per_step"], "build_us": g["build_ms"] * 1e3,
           "edges_per_s": E / (g["ms_per_step"] * 1e-3), "config_checks_per_s": g["config_checks_per_step"] / (g["ms_per_step"] * 1e-3),
           "valid_fraction": g["valid_fraction"], "valid_edges_per_s": g["valid_fraction"] * E / (g["ms_per_step"] * 1e-3),
           "nodes_per_step_graph": g["nodes_per_step_graph"], "min_step_ms": g["min_step_ms"], "max_step_ms": g["max_step_ms"],
           "run_to_run_spread": spread, "runs_ms_per_step": [r["ms_per_step"] for r in runs]}
    # the Python-driven loop for comparison (launch overhead visible, same edges every step)
    e_np = synth.random_edges(E)
    e = [cuda(x) for x in e_np]
    for i in range(3):
        w.build(clouds[i % 8]); w.check_edges(*e)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        w.build(clouds[i % len(clouds)])
        w.check_edges(*e)
    torch.cuda.synchronize()
    out["python_loop_ms_per_step"] = (time.perf_counter() - t0) / steps * 1e3
    if not args.no_cpu:
        out["cpu_baseline"] = cpu_edges(clouds_np[0], e_np, 2000)
        out["cpu_baseline"]["note"] = "edge validation only; the CPU rebuild (uniform-grid counting sort of the oracle) is not included"
    return out


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--only", default="")
    p.add_argument("--vpm", type=float, default=50.0)
    p.add_argument("--states", type=int, default=1024)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--edges", type=int, default=1_000_000)
    p.add_argument("--no-cpu", action="store_true")
    args = p.parse_args()
    assert torch.cuda.is_available(), "needs a CUDA device"
    for name, fn in (("C1", c1), ("C2", c2), ("C3", c3), ("C5", c5)):
        if args.only and name not in args.only.split(","):
            continue
        r = fn(args)
        r["voxels_per_metre"] = args.vpm
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
