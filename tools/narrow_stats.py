#!/usr/bin/env python
"""Work counters of the exact narrow phase on the bench workload (profiling aid).

Needs the -DTG_STATS build:  python -m tygrit_b200.build --stats
Run:  TYGRIT_B200_LIB=tygrit_b200/lib/libtygrit_b200_stats.so python tools/narrow_stats.py [--configs N] [--vpm V] [--cpm C]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tygrit_b200 import _native as N, synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--configs", type=int, default=1 << 22)
p.add_argument("--vpm", type=float, default=50.0)
p.add_argument("--cpm", type=float, default=0.0)
p.add_argument("--points", type=int, default=500_000)
a = p.parse_args()

w = CollisionWorld(device=0, max_points=1 << 20, voxels_per_metre=a.vpm, max_voxels=192 << 20)
if a.cpm > 0:
    w.set_cells_per_metre(a.cpm)
w.build(torch.from_numpy(synth.room_cloud(a.points)).cuda())
arm, base = synth.random_configs(a.configs, seed=42, base_box=(-3.5, 3.5, -2.5, 2.5))
arm_d, base_d = torch.from_numpy(arm).cuda(), torch.from_numpy(base).cuda()
out = (C.c_int64 * 16)()
N.check(N.lib().tg_debug_stats(out, 1), "tg_debug_stats")
w.counters()
valid = w.check_configs(arm_d, base_d)
torch.cuda.synchronize()
cnt = w.counters()
N.check(N.lib().tg_debug_stats(out, 1), "tg_debug_stats")
s = [int(v) for v in out]
items = max(cnt["narrow_items"], 1)
res = {
    "configs": a.configs, "vpm": a.vpm, "cpm": a.cpm, "items": cnt["narrow_items"], "items_per_config": cnt["narrow_items"] / a.configs,
    "spheres_listed_per_item": s[1] / items, "spheres_walked_per_item": s[2] / items,
    "row_batches_per_walked_sphere": s[3] / max(s[2], 1), "scan_steps_per_walked_sphere": s[4] / max(s[2], 1),
    "candidate_points_per_walked_sphere": s[9] / max(s[2], 1),
    "walked_small_R<0.13": s[5], "walked_mid": s[6], "walked_large_R>=0.2": s[7], "hits": s[8],
    "hit_fraction_of_items": s[8] / items, "free_fraction": float(valid.float().mean().item()),
}
print(json.dumps(res))
