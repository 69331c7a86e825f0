#!/usr/bin/env python
"""Localise the multi-rank slowdown of k_check_configs: per-launch phase-1 / phase-2 device times
of the bench step on THIS process's GPU under different process-group states.

  python tools/p1_time.py --tag alone                         one process, no torch.distributed
  torchrun --nproc-per-node 2 tools/p1_time.py --mode none    2 busy GPUs, no process group at all
  torchrun ... tools/p1_time.py --mode nccl                   NCCL communicator alive (one broadcast done)
  torchrun ... tools/p1_time.py --mode nccl-destroyed         communicator created, used, destroyed before timing
  NCCL_P2P_DISABLE=1 torchrun ... --mode nccl                 no peer mappings
Each rank prints one JSON line.
"""
import argparse, json, os, sys
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tygrit_b200 import synth  # noqa: E402
from tygrit_b200.engine import CollisionWorld  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--mode", default="none", choices=["none", "nccl", "nccl-destroyed", "gloo"])
p.add_argument("--touch", default="", help="comma list of NCCL uses before timing: blob (broadcast the structure into the env's blob), "
               "valid (all-gather the verdict tensor itself), staged (all-gather a separate staging copy of the verdicts)")
p.add_argument("--tag", default="")
p.add_argument("--device", type=int, default=-1)
p.add_argument("--steps", type=int, default=10)
p.add_argument("--vpm", type=float, default=72.0)
p.add_argument("--workload", default="uniform", choices=["uniform", "free"])
a = p.parse_args()
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
if a.device >= 0:
    local = a.device
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
import torch.distributed as dist
if a.mode != "none" and world > 1:
    if a.mode == "gloo":
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
        t = torch.ones(1 << 20, device=dev)
        dist.broadcast(t, 0)
        dist.all_reduce(t)
        torch.cuda.synchronize()
        if a.mode == "nccl-destroyed":
            dist.destroy_process_group()
w = CollisionWorld(device=local, max_points=1 << 20, voxels_per_metre=a.vpm, max_voxels=192 << 20)
w.build(torch.from_numpy(synth.room_cloud(500_000)).to(dev))
B = 64 << 20
arm, base = synth.device_configs(B, seed=42 + rank, device=dev)
if a.workload == "free":
    arm, base, _ = synth.free_base_configs(w, B, seed=42 + rank, device=dev)
valid = torch.empty(B, dtype=torch.uint8, device=dev)
for _ in range(3):
    w.check_configs_into(arm, base, valid)
torch.cuda.synchronize()
if dist.is_initialized() and a.touch:
    for what in a.touch.split(","):
        if what == "blob":
            n = torch.tensor([w.used_bytes() if rank == 0 else 0], dtype=torch.int64, device=dev)
            dist.broadcast(n, 0)
            dist.broadcast(w.blob_view(int(n.item())), 0)
            if rank != 0:
                w.adopt_blob(int(n.item()))
        elif what == "valid":
            out = torch.empty(world * B, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(out, valid)
        elif what == "staged":
            stage = valid.clone()
            out = torch.empty(world * B, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(out, stage)
    torch.cuda.synchronize()
if dist.is_initialized():
    dist.barrier()
w.set_timing(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    w.check_configs_into(arm, base, valid)
e1.record()
torch.cuda.synchronize()
tim = w.timing_read()
n = max(tim["launch_pairs"], 1)
print(json.dumps({"touch": a.touch, "tag": a.tag or a.mode, "rank": rank, "world": world, "device": local, "gpu": torch.cuda.get_device_name(dev),
                  "ms_per_step": e0.elapsed_time(e1) / a.steps, "phase1_ms_per_launch": tim["phase1_ms"] / n,
                  "phase2_ms_per_launch": tim["phase2_ms"] / n, "p2p_disable": os.environ.get("NCCL_P2P_DISABLE", "")}), flush=True)
if dist.is_initialized():
    dist.barrier()
    dist.destroy_process_group()
