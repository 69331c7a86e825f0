#!/usr/bin/env python
"""Randomised parity stress of the CUDA path against the fp64 C oracle (oracle/cpu_checker.c), beyond the fixed cases of
tests/: every round draws a cloud (room or tabletop, random size, random point radius), a grid resolution, a narrow-phase
cell shape and a batch size on either side of the coarse-variant threshold (2^18 configurations per call), then compares

  * verdicts of tg_check_configs              bit-exact outside |clearance| <= 1e-5 m,
  * verdicts of tg_check_configs_noalloc       identical to tg_check_configs (when the batch fits),
  * clearances of tg_check_configs             within 3e-6 m of the oracle (cap 0.25 m),
  * verdicts of tg_check_edges                 bit-exact outside the band, same sample counts.

  python tools/stress_parity.py [--rounds N] [--seed S]      (one GPU; prints one JSON line per round and a summary)
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cchecker  # noqa: E402  (checker only)
from tygrit_b200 import _native as N, synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--rounds", type=int, default=12)
p.add_argument("--seed", type=int, default=0)
a = p.parse_args()
dev = torch.device("cuda", 0)
BAND = 1e-5
fails = 0
t_all = time.time()
for rnd in range(a.rounds):
    rng = np.random.default_rng(1000 * a.seed + rnd)
    kind = ["room", "tabletop"][int(rng.integers(0, 2))]
    n_pts = int(rng.integers(2_000, 400_000))
    cloud = synth.room_cloud(n_pts, seed=int(rng.integers(1, 1 << 30))) if kind == "room" else synth.tabletop_cloud(n_pts, seed=int(rng.integers(1, 1 << 30)))
    r_pt = float(rng.choice([0.0, 0.01, 0.03, 0.05]))
    vpm = float(rng.choice([30.0, 50.0, 64.0, 72.0, 90.0]))
    cpmx, cpm = float(rng.choice([32.0, 64.0, 96.0])), float(rng.choice([10.0, 14.0, 20.0]))
    B = int(rng.choice([1, 31, 4097, 100_003, (1 << 18) - 1, 1 << 18, 300_001]))
    box = (-3.5, 3.5, -2.5, 2.5) if kind == "room" else (-1.2, 1.2, -1.2, 1.2)
    w = CollisionWorld(device=0, max_points=1 << 19, voxels_per_metre=vpm, max_voxels=192 << 20, point_radius=r_pt)
    w.set_cell_shape(cpmx, cpm)
    w.build(torch.from_numpy(cloud).to(dev))
    arm, base = synth.random_configs(B, seed=int(rng.integers(1, 1 << 30)), base_box=box)
    ta, tb = torch.from_numpy(arm).to(dev), torch.from_numpy(base).to(dev)
    cc = cchecker.CCloud(cloud, point_radius=r_pt)
    n_or = min(B, 120_000)
    cl = cc.clearance(arm[:n_or], base[:n_or], cap=0.25)
    valid, clr = w.check_configs(ta, tb, clearance=True)
    v_only = w.check_configs(ta, tb)
    got = valid[:n_or].cpu().numpy().astype(bool)
    dec = np.abs(cl) > BAND
    bad_v = int(np.count_nonzero(got[dec] != (cl > 0)[dec]))
    same = bool(torch.equal(valid, v_only))
    err = np.abs(clr[:n_or].cpu().numpy() - cl)
    bad_c = int(np.count_nonzero(err > 3e-6))
    na = None
    if B <= N.TG_NOALLOC_MAX_CONFIGS:
        out = torch.full((B,), 9, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            N.check(N.lib().tg_check_configs_noalloc(w._h, ta.data_ptr(), tb.data_ptr(), B, out.data_ptr(),
                                                     torch.cuda.current_stream(dev).cuda_stream), "tg_check_configs_noalloc")
        na = bool(torch.equal(out, v_only))
    # obstacle + attachment spheres (the lane-per-configuration clearance kernel), self-collision term on every other round
    from oracle import collide  # noqa: E402  (numpy + cKDTree fp64 oracle; knows obstacles and attachments)
    n_x = min(B, 3000)
    n_obs = int(rng.integers(0, 5))
    obs = np.concatenate([rng.uniform([box[0], box[2], 0.1], [box[1], box[3], 1.8], size=(n_obs, 3)), rng.uniform(0.03, 0.3, size=(n_obs, 1))], axis=1).astype(np.float32)
    n_att = int(rng.integers(0, 4))
    att = {"spheres": np.concatenate([rng.uniform(-0.05, 0.25, size=(n_att, 3)), rng.uniform(0.02, 0.08, size=(n_att, 1))], axis=1),
           "pos": tuple(rng.uniform(-0.03, 0.03, size=3)), "quat": tuple(q / np.linalg.norm(q) for q in [rng.normal(size=4)])[0]}
    self_on = bool(rnd & 1)
    w.set_obstacles(obs)
    w.set_attachment(att["spheres"], att["pos"], att["quat"])
    w.set_self_collision(self_on)
    xv, xc = w.check_configs(ta[:n_x], tb[:n_x], clearance=True)
    xcl = collide.CloudOracle(cloud, point_radius=r_pt, obstacles=obs).config_clearance(arm[:n_x], base[:n_x], attachment=att if n_att else None)
    if self_on:
        xcl = np.minimum(xcl, cchecker.self_clearance(arm[:n_x]))
    xdec = np.abs(xcl) > BAND
    bad_xv = int(np.count_nonzero(xv.cpu().numpy().astype(bool)[xdec] != (xcl > 0)[xdec]))
    xerr = np.abs(xc.cpu().numpy() - np.minimum(xcl, 0.25))
    bad_xc = int(np.count_nonzero(xerr > 3e-6))
    w.set_obstacles(np.empty((0, 4), np.float32))
    w.set_attachment(np.empty((0, 4), np.float32))
    w.set_self_collision(False)
    # edges
    E = int(rng.choice([1, 77, 5_000, 60_000]))
    a0, b0, a1, b1 = synth.random_edges(E, seed=int(rng.integers(1, 1 << 30)), base_box=box, rrt_like=bool(rng.integers(0, 2)))
    ok, nchk = w.check_edges_host(a0, b0, a1, b1, 8, 32.0)
    n_e = min(E, 3_000)
    ecl, ens = cc.edge_clearance(a0[:n_e], b0[:n_e], a1[:n_e], b1[:n_e], cap=1e-3)
    edec = np.abs(ecl) > BAND
    bad_e = int(np.count_nonzero(ok[:n_e].astype(bool)[edec] != (ecl > 0)[edec]))
    rec = {"round": rnd, "cloud": kind, "points": n_pts, "r_pt": r_pt, "vpm": vpm, "cells": [cpmx, cpm], "configs": B, "edges": E,
           "free_fraction": float(got.mean()), "verdict_mismatches": bad_v, "clearance_over_3e-6": bad_c, "max_clearance_err": float(err.max()),
           "with_and_without_clearance_equal": same, "noalloc_equal": na, "edge_mismatches": bad_e, "edge_valid_fraction": float(ok.mean()),
           "extras": {"obstacles": n_obs, "attached": n_att, "self": self_on, "verdict_mismatches": bad_xv, "clearance_over_3e-6": bad_xc,
                      "max_clearance_err": float(xerr.max())}}
    print(json.dumps(rec), flush=True)
    if bad_v or bad_c or bad_e or bad_xv or bad_xc or not same or na is False:
        fails += 1
    w.close()
print(json.dumps({"rounds": a.rounds, "failed_rounds": fails, "seconds": time.time() - t_all}))
sys.exit(1 if fails else 0)
