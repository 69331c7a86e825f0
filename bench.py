#!/usr/bin/env python
"""bench.py — Fetch configurations collision-checked per second against a point cloud.

Metric (BASELINE.json): "Fetch configs collision-checked/sec vs point cloud, 1/2/4/8 B200 vs VAMP CPU".
Workload = BASELINE config[3] (C4): a bulk check of 64M random Fetch configurations per GPU against
the 500k-point synthetic room cloud; the cloud structure is built on rank 0 and NCCL-broadcast.
One "step" = one pass of the hot path (FK + grid classification + exact narrow phase) over the batch.

  python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
  python bench.py --impl reference ...                     CPU arm: the oracle port on all host threads

Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fetch_configs_collision_checked_per_sec"
UNIT = "configs/s"
F_CFG = 4620.0          # algorithmic flop per configuration (FK 2400 + 111 probes x 20)
B_CFG = 49.0            # algorithmic HBM bytes per configuration (44 in + 1 out + 4 spare)
CLOUD_POINTS = 500_000
BASE_BOX = (-3.5, 3.5, -2.5, 2.5)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--configs", type=int, default=64 * 1024 * 1024, help="configurations per GPU per step")
    p.add_argument("--cpu-sample", type=int, default=4_000_000, help="configurations per CPU-baseline sample")
    p.add_argument("--vpm", type=float, default=72.0, help="distance-grid voxels per metre (static cloud: the finer grid pays, sweep in profiles/)")
    p.add_argument("--cpm", type=float, default=0.0, help="narrow-phase cells per metre (0 = library default)")
    p.add_argument("--cpmx", type=float, default=0.0, help="narrow-phase cells per metre along x (with --cpm = y/z; 0 = same as --cpm)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-cpu", action="store_true")
    return p.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


def traffic_from_profile(cfg_per_launch, vpm):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_check_configs launch from the committed
    `ncu --set full` capture (profiles/traffic.json); None when the capture was taken at another launch shape."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    t = json.load(open(path)).get("k_check_configs")
    if not t or int(t["configs_per_launch"]) != int(cfg_per_launch) or float(t.get("voxels_per_metre", vpm)) != float(vpm):
        return None
    return {"bytes_per_launch": t["dram_bytes_read"] + t["dram_bytes_write"], "algorithmic_bytes_per_launch": B_CFG * cfg_per_launch,
            "source": t["source"]}


_REAL_STDOUT = 1


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {
                "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            }
            while not self.stop_flag:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.004)  # the timed region is ~40 ms: several samples inside it
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_arm(args, cloud, n_steps, n_warm):
    """The reference arm / cpu_baseline: the oracle's C port of the verdict arithmetic, all host threads."""
    from oracle import cchecker
    from tygrit_b200 import synth

    cc = cchecker.CCloud(cloud)
    threads = cchecker.threads()
    arm, base = synth.random_configs(args.cpu_sample, seed=42, base_box=BASE_BOX)
    for _ in range(min(n_warm, 1)):
        cc.check_configs_f32(arm[:20_000], base[:20_000])
    t0 = time.perf_counter()
    free = 0
    for _ in range(n_steps):
        free += int(cc.check_configs_f32(arm, base).sum())
    dt = time.perf_counter() - t0
    return {
        "value": args.cpu_sample * n_steps / dt, "unit": UNIT, "cores": threads, "kind": "port",
        "sample": f"{n_steps} x {args.cpu_sample} configs (seed 42) vs the same {len(cloud)}-point room cloud, "
                  f"fp32 uniform-grid checker oracle/cpu_checker.c, OpenMP {threads} threads; VAMP itself unavailable on this box",
        "ms_per_step": dt / n_steps * 1e3, "free_fraction": free / (args.cpu_sample * n_steps),
    }


def emit(line: dict) -> None:
    """The ONE JSON line of the contract, written to the process's original stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def main():
    # Libraries chat on stdout (NCCL prints "NCCL version ..." when its communicator comes up): everything except the
    # JSON line goes to stderr.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from tygrit_b200 import synth

    config = {
        "workload": f"C4: bulk check of {args.configs} random Fetch configs per GPU vs {CLOUD_POINTS}-point synthetic room cloud",
        "configs_per_gpu_per_step": args.configs, "cloud_points": CLOUD_POINTS, "point_radius": 0.03,
        "voxels_per_metre": args.vpm, "l2": "inputs (44 B/config = %.1f GB/step) are larger than L2" % (args.configs * 44 / 1e9),
        "parallelism": f"configs sharded over {world} rank(s), cloud structure broadcast",
    }

    if args.impl == "reference":
        if rank != 0:
            return
        cloud = synth.room_cloud(CLOUD_POINTS)
        r = cpu_arm(args, cloud, max(1, args.steps), args.warmup)
        line = {
            "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        emit(line)
        return

    import torch
    import torch.distributed as dist

    from tygrit_b200.engine import CollisionWorld

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- world: rank 0 ingests the cloud and builds; the structure is broadcast ------------------
    w = CollisionWorld(device=local, max_points=1 << 20, voxels_per_metre=args.vpm, max_voxels=192 << 20)
    if args.cpm > 0 and args.cpmx > 0:
        w.set_cell_shape(args.cpmx, args.cpm)
    elif args.cpm > 0:
        w.set_cells_per_metre(args.cpm)
    t_build = t_bcast = 0.0
    cloud = None
    if rank == 0:
        cloud = synth.room_cloud(CLOUD_POINTS)
        cd = torch.from_numpy(cloud).to(dev)
        w.build(cd)                                    # warm-up build
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        w.build(cd)
        e1.record()
        torch.cuda.synchronize()
        t_build = e0.elapsed_time(e1)
    if world > 1:
        nbytes = torch.tensor([w.used_bytes() if rank == 0 else 0], dtype=torch.int64, device=dev)
        dist.broadcast(nbytes, 0)
        view = w.blob_view(int(nbytes.item()))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dist.broadcast(view, 0)
        torch.cuda.synchronize()
        t_bcast = (time.perf_counter() - t0) * 1e3
        if rank != 0:
            w.adopt_blob(int(nbytes.item()))
    stats = w.stats()

    # ---- inputs: generated on the device (seeded per rank), resident in HBM ------------------------
    B = args.configs
    gen = torch.Generator(device=dev)
    gen.manual_seed(42 + rank)
    from tygrit_b200.fetch_model import JOINT_LOWER, JOINT_UPPER

    lo = torch.tensor(JOINT_LOWER, dtype=torch.float32, device=dev)
    hi = torch.tensor(JOINT_UPPER, dtype=torch.float32, device=dev)
    arm = torch.rand(B, 8, generator=gen, device=dev) * (hi - lo) + lo
    blo = torch.tensor([BASE_BOX[0], BASE_BOX[2], -np.pi], dtype=torch.float32, device=dev)
    bhi = torch.tensor([BASE_BOX[1], BASE_BOX[3], np.pi], dtype=torch.float32, device=dev)
    base = torch.rand(B, 3, generator=gen, device=dev) * (bhi - blo) + blo
    valid = torch.empty(B, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        w.check_configs_into(arm, base, valid)
    barrier()
    w.counters()
    w.set_timing(True)
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        w.check_configs_into(arm, base, valid)
    e1.record()
    barrier()
    elapsed_ms = e0.elapsed_time(e1)
    tim = w.timing_read()
    w.set_timing(False)
    cnt = w.counters()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    free_frac = float(valid.float().mean().item())
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())
    ms_per_step = elapsed_ms / args.steps
    value = world * B / (ms_per_step * 1e-3)

    # ---- e2e: the reference-facing call with HOST buffers (pinned), copies inside the timed region ---
    e2e = None
    if not args.no_e2e:
        h_arm = torch.empty(B, 8, dtype=torch.float32, pin_memory=True)
        h_base = torch.empty(B, 3, dtype=torch.float32, pin_memory=True)
        h_valid = torch.empty(B, dtype=torch.uint8, pin_memory=True)
        h_arm.copy_(arm)
        h_base.copy_(base)
        torch.cuda.synchronize()
        a_np, b_np, v_np = h_arm.numpy(), h_base.numpy(), h_valid.numpy()
        n_e2e = max(2, min(args.steps, 5))
        w.check_configs_host(a_np, b_np, out=v_np)      # warm-up (allocates the pipeline buffers)
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            w.check_configs_host(a_np, b_np, out=v_np)
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        assert bool((torch.from_numpy(v_np[: 1 << 20]).to(dev) == valid[: 1 << 20]).all()), "host path disagrees with device path"
        e2e = {"value": world * B * n_e2e / dt, "unit": UNIT, "h2d_bytes_per_step": B * 44, "d2h_bytes_per_step": B,
               "steps": n_e2e, "ms_per_step": dt / n_e2e * 1e3, "api": "CollisionWorld.check_configs_host -> tg_check_configs_host"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    props = torch.cuda.get_device_properties(dev)
    fp32_peak = props.multi_processor_count * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12      # TFLOP/s, non-tensor
    k1_ms = tim["phase1_ms"] / max(tim["launch_pairs"], 1)                                  # avg k_check_configs launch
    cfg_per_launch = B * args.steps / max(tim["launch_pairs"], 1)
    achieved = cfg_per_launch * F_CFG / (k1_ms * 1e-3) / 1e12
    roofline = {
        "kernel": "k_check_configs (fused FK + distance-grid classification)", "bound": "fp32",
        "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
        "peak_source": f"{props.multi_processor_count} SMs x 128 lanes x 2 flop x {pk['sm_max_mhz']:.0f} MHz ({pk['source']} sm_max_mhz)",
        "algorithmic_flop_per_config": F_CFG, "configs_per_launch": cfg_per_launch, "avg_launch_ms": k1_ms,
        "share_of_step": tim["phase1_ms"] / elapsed_ms if world == 1 else None,
        "phase2_share_of_step": tim["phase2_ms"] / elapsed_ms if world == 1 else None,
        "traffic": traffic_from_profile(cfg_per_launch, args.vpm),
        "hbm": {"achieved": cfg_per_launch * B_CFG / (k1_ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": cfg_per_launch * B_CFG / (k1_ms * 1e-3) / 1e9 / pk["hbm_gbs"], "algorithmic_bytes_per_config": B_CFG,
                "peak_source": pk["source"]},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": config, "e2e": e2e, "gpu_launches": cnt["launches"], "clocks": sampler.result(),
        "roofline": roofline,
        "world": {"build_ms": t_build, "broadcast_ms": t_bcast, "grid": [stats["nx"], stats["ny"], stats["nz"]],
                  "slack_m": stats["slack"], "free_fraction": free_frac,
                  "narrow_items_per_config": cnt["narrow_items"] / (B * args.steps),
                  "definite_collision_fraction": cnt["definite_collisions"] / (B * args.steps),
                  "overflow_inline": cnt["overflow_inline"]},
    }
    if world == 1 and not args.no_cpu:
        r = cpu_arm(args, cloud, 1, 1)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
