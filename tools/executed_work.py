#!/usr/bin/env python
"""Executed-work counters of the configuration check on the bench workload ("probes per config and
narrow-phase tests_per_config from a kernel counter build").

Needs the -DTG_STATS build (``python -m tygrit_b200.build --stats``; ``__graft_entry__.build()`` makes it) and is run
by ``bench.py`` in a subprocess with ``TYGRIT_B200_LIB`` pointing at it — the counting build is never the timed one.

  TYGRIT_B200_LIB=tygrit_b200/lib/libtygrit_b200_stats.so python tools/executed_work.py [--configs N] [--vpm V] [--workload free]

Prints one JSON object: distance-grid probes per configuration (executed and skipped by a derived bound), ambiguous
items, narrow-phase spheres walked and candidate points tested per configuration.
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tygrit_b200 import _native as N, synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--configs", type=int, default=1 << 22)
p.add_argument("--vpm", type=float, default=100.0)
p.add_argument("--points", type=int, default=500_000)
p.add_argument("--workload", default="uniform", choices=["uniform", "free"])
a = p.parse_args()

if "stats" not in os.path.basename(N.LIB_PATH):
    print(json.dumps({"error": "TYGRIT_B200_LIB does not point at the -DTG_STATS build"}))
    sys.exit(0)
w = CollisionWorld(device=0, max_points=1 << 20, voxels_per_metre=a.vpm, max_voxels=192 << 20)
w.build(torch.from_numpy(synth.room_cloud(a.points)).cuda())
dev = torch.device("cuda", 0)
arm_d, base_d = synth.device_configs(a.configs, seed=42, device=dev)
if a.workload == "free":
    arm_d, base_d, _ = synth.free_base_configs(w, a.configs, seed=42, device=dev)
out = (C.c_int64 * 16)()
N.check(N.lib().tg_debug_stats(out, 1), "tg_debug_stats")
w.counters()
valid = w.check_configs(arm_d, base_d)
torch.cuda.synchronize()
cnt = w.counters()
N.check(N.lib().tg_debug_stats(out, 1), "tg_debug_stats")
s = [int(v) for v in out]
B = a.configs
print(json.dumps({
    "configs": B, "voxels_per_metre": a.vpm, "workload": a.workload, "free_fraction": float(valid.float().mean().item()),
    "coarse_lookups_per_config": s[13] / B, "probes_per_config": s[10] / B, "stage_b_fraction": s[12] / B, "probes_skipped_by_derived_bound_per_config": s[11] / B,
    "narrow_items_per_config": cnt["narrow_items"] / B, "definite_collision_fraction": cnt["definite_collisions"] / B,
    "narrow_spheres_listed_per_config": s[1] / B, "narrow_spheres_walked_per_config": s[2] / B,
    "narrow_point_tests_per_config": s[9] / B, "narrow_row_batches_per_config": s[3] / B,
}))
