#!/usr/bin/env python
"""Narrow-phase cell shape (tg_env_set_cell_shape) against the two small workloads it matters for:
C1 (4096 configurations vs the dense 50k-point tabletop cloud, one-call latency) and C2 (100k edges vs the 200k-point room cloud).
usage: python tools/bench_cellshape.py  (one GPU; one line per shape, x / yz cells per metre)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.bench_configs import cuda, timed  # noqa: E402
from tygrit_b200 import synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

SHAPES = [(20, 20), (40, 10), (40, 14), (40, 20), (60, 14), (60, 20), (80, 20), (80, 14), (30, 14)]


def main():
    top, room = cuda(synth.tabletop_cloud(50_000)), cuda(synth.room_cloud(200_000))
    arm, base = synth.random_configs(4096, base_box=(-1, 1, -1, 1))
    a, b = cuda(arm), cuda(base)
    v = torch.empty(4096, dtype=torch.uint8, device="cuda")
    e = [cuda(x) for x in synth.random_edges(100_000)]
    for cx, cyz in SHAPES:
        w = CollisionWorld(max_points=1 << 20, voxels_per_metre=50.0)
        w.set_cell_shape(cx, cyz)
        w.build(top)
        c1 = timed(lambda: w.check_configs_into(a, b, v), 200, warm=20) * 1e3
        w.build(room)
        c2 = timed(lambda: w.check_edges(*e), 100, warm=50)
        print(f"x {cx:5.1f} yz {cyz:5.1f} cells/m   C1 {c1:7.2f} us   C2 {c2:7.4f} ms", flush=True)
        w.close()


if __name__ == "__main__":
    main()
