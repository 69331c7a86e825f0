#!/usr/bin/env python
"""bench.py — Fetch configurations collision-checked per second against a point cloud.

Metric (BASELINE.json): "Fetch configs collision-checked/sec vs point cloud, 1/2/4/8 B200 vs VAMP CPU".
Workload = BASELINE config[3] (C4): a bulk check of 64 Mi random Fetch configurations against the 500k-point
synthetic room cloud.  One "step" = one pass of the hot path (FK + grid classification + exact narrow phase,
verdict bytes written) over one batch.

  python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
  python bench.py --impl reference ...                     CPU arm: the oracle's C port on all host threads

N = 1: the step is the check of 64 Mi device-resident configurations.
N > 1: the step is the whole sharded job — NCCL broadcast of the cloud structure from rank 0, the check of every
rank's block, and the all-gather of the verdicts, packed to bits (one gather per launch pair) — all inside the timed region.  `value` is the weak-scaling figure (64 Mi configurations per GPU);
the same line carries `strong` (BASELINE's C4 as stated: 64 Mi configurations in total, split over the N ranks)
and `replicas_only` (the step without any collective, what round 1 timed).

Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fetch_configs_collision_checked_per_sec"
UNIT = "configs/s"
F_CFG = 4620.0          # algorithmic flop per configuration (FK 2400 + 111 probes x 20)
F_TEST = 9.0            # flop per executed narrow-phase point test
B_CFG = 49.0            # algorithmic HBM bytes per configuration (44 in + 1 out + 4 clearance)
CLOUD_POINTS = 500_000
BASE_BOX = (-3.5, 3.5, -2.5, 2.5)
LAUNCH_CONFIGS = 64 * 1024 * 1024   # configurations per launch pair of the device API (8 x the ambiguous queue)
PARITY_SAMPLE = 1 << 20


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--configs", type=int, default=64 * 1024 * 1024, help="configurations per GPU per step (weak) = total of the strong run")
    p.add_argument("--cpu-sample", type=int, default=8_000_000, help="configurations per CPU-baseline sample")
    p.add_argument("--vpm", type=float, default=100.0,
                   help="distance-grid voxels per metre (the library's maximum): the cloud of C4 is static, and since phase 1 takes most of its "
                        "proofs from the coarse shared-memory grid its time no longer depends on the fine grid's size, while the narrow phase "
                        "shrinks with the voxel (sweeps in profiles/r02g_vpm_sweep.txt and r03s_vpm_sweep.txt); the library default stays 50")
    p.add_argument("--cpm", type=float, default=0.0, help="narrow-phase cells per metre (0 = library default)")
    p.add_argument("--cpmx", type=float, default=0.0, help="narrow-phase cells per metre along x (with --cpm = y/z; 0 = same as --cpm)")
    p.add_argument("--workload", default="uniform", choices=["uniform", "free"],
                   help="headline batch: 'uniform' = base uniform over the floor (C4 as stated); 'free' = base on free floor (planner-like)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-cpu", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the free-workload, clearance, strong-scaling and executed-work legs")
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
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis:  # NVML enumerates physical devices
                try:
                    idx = int(vis.split(",")[self.index])
                except Exception:
                    pass
            h = nv.nvmlDeviceGetHandleByIndex(idx)
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
                time.sleep(0.004)  # the timed region is tens of ms: several samples inside it
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_arm(args, cloud, n_steps, n_warm):
    """The reference arm / cpu_baseline: the oracle's C port of the verdict arithmetic (coarse-to-fine scalar fp32
    checker), on ALL host threads — the count is set explicitly, torchrun's OMP_NUM_THREADS=1 is not inherited —
    and, on a smaller sample, on ONE thread (VAMP is single-threaded per query: the like-for-like figure)."""
    from oracle import cchecker
    from tygrit_b200 import synth

    cc = cchecker.CCloud(cloud)
    threads = cchecker.host_threads()
    arm, base = synth.random_configs(args.cpu_sample, seed=42, base_box=BASE_BOX)
    cc.check_configs_f32(arm[:100_000], base[:100_000], threads=threads)
    t0 = time.perf_counter()
    free = 0
    for _ in range(n_steps):
        free += int(cc.check_configs_f32(arm, base, threads=threads).sum())
    dt = time.perf_counter() - t0
    n1 = max(100_000, args.cpu_sample // 16)
    t1 = time.perf_counter()
    cc.check_configs_f32(arm[:n1], base[:n1], threads=1)
    dt1 = time.perf_counter() - t1
    cchecker.set_threads(threads)
    return {
        "value": args.cpu_sample * n_steps / dt, "unit": UNIT, "cores": threads, "kind": "port",
        "sample": f"{n_steps} x {args.cpu_sample} configs (numpy seed 42, same distribution as the GPU batch) vs the same {len(cloud)}-point "
                  f"room cloud; fp32 coarse-to-fine uniform-grid checker oracle/cpu_checker.c (link by link, enclosing sphere first, "
                  f"early exit), OpenMP {threads} threads; VAMP itself unavailable on this box",
        "ms_per_step": dt / n_steps * 1e3, "free_fraction": free / (args.cpu_sample * n_steps),
        "one_core": {"value": n1 / dt1, "unit": UNIT, "cores": 1, "sample": f"{n1} configs of the same batch"},
    }


def emit(line: dict) -> None:
    """The ONE JSON line of the contract, written to the process's original stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def executed_work(args, workload):
    """Probes and narrow-phase point tests per configuration from the -DTG_STATS counting build (a separate
    process on a 4 Mi-configuration sample of the same distribution; never the timed library)."""
    lib = os.path.join(ROOT, "tygrit_b200", "lib", "libtygrit_b200_stats.so")
    if not os.path.exists(lib):
        return None
    env = dict(os.environ, TYGRIT_B200_LIB=lib)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "executed_work.py"), "--vpm", str(args.vpm),
                            "--workload", workload], env=env, capture_output=True, text=True, timeout=300)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # pragma: no cover
        return {"error": f"{type(e).__name__}: {e}"}


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
        "workload": f"C4: bulk check of {args.configs} random Fetch configs per GPU vs {CLOUD_POINTS}-point synthetic room cloud"
                    + (" (base on free floor)" if args.workload == "free" else ""),
        "configs_per_gpu_per_step": args.configs, "cloud_points": CLOUD_POINTS, "point_radius": 0.03,
        "voxels_per_metre": args.vpm, "l2": "inputs (44 B/config = %.1f GB/step) are larger than L2" % (args.configs * 44 / 1e9),
        "parallelism": f"configs sharded over {world} rank(s); per step: structure NCCL-broadcast from rank 0, verdict bits all-gathered"
                       if world > 1 else "1 rank",
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
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "one_core")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        emit(line)
        return

    import torch
    import torch.distributed as dist

    from tygrit_b200 import engine
    from tygrit_b200.engine import CollisionWorld
    from tygrit_b200.shard import shard_range

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
    t_build = 0.0
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
    blob_bytes = 0
    if world > 1:
        nbytes = torch.tensor([w.used_bytes() if rank == 0 else 0], dtype=torch.int64, device=dev)
        dist.broadcast(nbytes, 0)
        blob_bytes = int(nbytes.item())
        view = w.blob_view(blob_bytes)
        dist.broadcast(view, 0)
        if rank != 0:
            w.adopt_blob(blob_bytes)
        torch.cuda.synchronize()
    stats = w.stats()

    # ---- inputs: generated on the device (seeded per rank), resident in HBM ------------------------
    B = args.configs
    floor_free = None
    if args.workload == "free":
        arm, base, floor_free = synth.free_base_configs(w, B, seed=42 + rank, device=dev, base_box=BASE_BOX)
    else:
        arm, base = synth.device_configs(B, seed=42 + rank, device=dev, base_box=BASE_BOX)
    valid = torch.empty(B, dtype=torch.uint8, device=dev)
    # the gather moves verdict BITS (tg_pack_bits: 1/8 of the bytes): [chunk][rank][config / 8]
    packed = torch.empty((B + 7) // 8 + 8, dtype=torch.uint8, device=dev) if world > 1 else None
    gathered = torch.empty(world * ((B + 7) // 8 + 8), dtype=torch.uint8, device=dev) if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def job(arm_t, base_t, valid_t, collectives: bool):
        """One sharded job on this rank's block: [broadcast structure] -> check -> [pack + all-gather the verdict bits]."""
        n = arm_t.shape[0]
        if not (collectives and world > 1):
            w.check_configs_into(arm_t, base_t, valid_t)
            return
        dist.broadcast(view, 0)                   # NCCL: rank 0's blob -> every rank's own blob storage
        if rank != 0:
            w.adopt_blob(blob_bytes)              # rebase pointers, refresh the texture copy of the distance grid
        for off in range(0, n, LAUNCH_CONFIGS):   # one launch pair per chunk, then the gather of its verdict bits
            m = min(LAUNCH_CONFIGS, n - off)
            w.check_configs_into(arm_t[off:off + m], base_t[off:off + m], valid_t[off:off + m])
            o8, m8 = off // 8, (m + 7) // 8
            engine.pack_verdicts(valid_t[off:off + m], packed[o8:o8 + m8])
            # 4 MiB per rank and launch pair: 35 us alone.  Not overlapped with the next pair's kernels on purpose: an NCCL kernel
            # that shares the SMs with the persistent, whole-SM blocks of phase 1 stretched phase 1 by 0.4 ms (r02u)
            dist.all_gather_into_tensor(gathered[world * o8: world * (o8 + m8)], packed[o8:o8 + m8])

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        per_rank = [ms]
        if world > 1:
            allms = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allms, t)
            per_rank = [float(x.item()) for x in allms]
        return max(per_rank) / steps, [x / steps for x in per_rank]

    W = max(args.warmup, 3)
    for _ in range(W):
        job(arm, base, valid, True)
    barrier()
    w.counters()
    w.set_timing(True)
    sampler = ClockSampler(local)
    sampler.start()
    ms_per_step, ms_per_rank = timed(lambda: job(arm, base, valid, True), args.steps)
    tim = w.timing_read()
    w.set_timing(False)
    cnt = w.counters()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.result()
    free_frac = float(valid.float().mean().item())
    value = world * B / (ms_per_step * 1e-3)

    # ---- multi-GPU extras: the step without collectives, and C4 as stated (strong scaling) ----------
    multi = None
    if world > 1:
        rep_ms, _ = timed(lambda: job(arm, base, valid, False), max(3, args.steps // 2))
        # exposed cost of each collective, measured alone on the same buffers
        bc_ms, _ = timed(lambda: dist.broadcast(view, 0), 10)
        m0 = min(B, LAUNCH_CONFIGS)
        ag_ms, _ = timed(lambda: dist.all_gather_into_tensor(gathered[: world * (m0 // 8)], packed[: m0 // 8]), 10)
        # the gathered bits of the last job are the verdict bytes: rank r's block, unpacked, equals what rank r computed
        own = engine.unpack_verdicts(gathered[rank * (m0 // 8): (rank + 1) * (m0 // 8)], m0)
        assert bool((own == valid[:m0]).all()), "gathered verdict bits differ from the verdict bytes"
        lo, hi = shard_range(B, rank, world)
        s_arm, s_base, s_valid = arm[: hi - lo], base[: hi - lo], valid[: hi - lo]   # this rank's block of the 64 Mi job
        for _ in range(3):
            job(s_arm, s_base, s_valid, True)
        st_ms, st_rank = timed(lambda: job(s_arm, s_base, s_valid, True), args.steps)
        st_rep_ms, _ = timed(lambda: job(s_arm, s_base, s_valid, False), args.steps)
        all_clocks = [None] * world
        dist.all_gather_object(all_clocks, clocks)
        multi = {
            "replicas_only": {"ms_per_step": rep_ms, "value": world * B / (rep_ms * 1e-3),
                              "note": "no collective inside the timed region (what round 1 reported)"},
            "collectives_alone": {"broadcast_ms": bc_ms, "broadcast_bytes": blob_bytes, "broadcast_GBps": blob_bytes / (bc_ms * 1e-3) / 1e9,
                                  "all_gather_ms": ag_ms, "all_gather_bytes_per_rank": m0 // 8,
                                  "all_gather_GBps_in": (world - 1) * (m0 // 8) / (ag_ms * 1e-3) / 1e9,
                                  "all_gather_payload": "verdict bits (tg_pack_bits), 1 bit per configuration"},
            "ms_per_step_per_rank": ms_per_rank, "clocks_per_rank": all_clocks,
            "strong": {"scaling": "strong", "configs_total": B, "configs_per_gpu": hi - lo, "ms_per_step": st_ms, "value": B / (st_ms * 1e-3),
                       "ms_per_step_per_rank": st_rank, "ms_per_step_without_collectives": st_rep_ms,
                       "note": "BASELINE configs[3] as stated: 64 Mi configurations in total split over the ranks; step = broadcast + check + all-gather"},
        }

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
        del h_arm, h_base, h_valid

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- parity sample: the benchmarked configurations themselves against the fp64 oracle -----------
    from oracle import cchecker

    cchecker.set_threads(cchecker.host_threads())
    ns = min(PARITY_SAMPLE, B)
    s_arm, s_base, s_val = arm[:ns].cpu().numpy(), base[:ns].cpu().numpy(), valid[:ns].cpu().numpy().astype(bool)
    t0 = time.perf_counter()
    cl = cchecker.CCloud(cloud).clearance(s_arm, s_base, cap=1e-3)
    decisive = np.abs(cl) > 1e-5
    mism = int(np.count_nonzero(s_val[decisive] != (cl > 0)[decisive]))
    parity = {"n": ns, "mismatches": mism, "in_band": int(ns - np.count_nonzero(decisive)), "band_m": 1e-5,
              "oracle": "oracle/cpu_checker.c tgo_clearance_f64 (fp64 exact clearance) on the first n GPU-generated configurations of the timed batch",
              "oracle_free_fraction": float((cl > 0).mean()), "seconds": time.perf_counter() - t0}
    assert mism == 0, f"bench batch: {mism} verdicts differ from the fp64 oracle outside the 1e-5 m band"

    extras = {}
    if world == 1 and not args.no_extras:
        # free-heavy (planner-like) batch on the same world: base on free floor, arm uniform
        f_arm, f_base, f_floor = (arm, base, floor_free) if args.workload == "free" else synth.free_base_configs(w, B, 42, dev, BASE_BOX)
        f_valid = torch.empty(B, dtype=torch.uint8, device=dev)
        for _ in range(2):
            w.check_configs_into(f_arm, f_base, f_valid)
        w.counters()
        f_ms, _ = timed(lambda: w.check_configs_into(f_arm, f_base, f_valid), max(3, args.steps // 2))
        f_cnt = w.counters()
        nf = min(1 << 18, B)
        f_cl = cchecker.CCloud(cloud).clearance(f_arm[:nf].cpu().numpy(), f_base[:nf].cpu().numpy(), cap=1e-3)
        f_dec = np.abs(f_cl) > 1e-5
        f_mism = int(np.count_nonzero(f_valid[:nf].cpu().numpy().astype(bool)[f_dec] != (f_cl > 0)[f_dec]))
        assert f_mism == 0, f"free-workload batch: {f_mism} verdicts differ from the fp64 oracle"
        steps_f = max(3, args.steps // 2)
        extras["free_workload"] = {
            "value": B / (f_ms * 1e-3), "unit": UNIT, "ms_per_step": f_ms, "free_fraction": float(f_valid.float().mean().item()),
            "floor_fraction_free_for_tucked_robot": f_floor, "narrow_items_per_config": f_cnt["narrow_items"] / (B * steps_f),
            "parity_sample": {"n": nf, "mismatches": f_mism, "in_band": int(nf - np.count_nonzero(f_dec))},
            "workload": "base uniform over the floor positions where the tucked robot is free, arm uniform in the joint limits",
            "executed": executed_work(args, "free"),
        }
        # verdict + clearance in one call (north_star: "per-config verdicts and clearances")
        nc = min(B, 8 * 1024 * 1024)
        c_valid, c_clear = w.check_configs(arm[:nc], base[:nc], clearance=True)
        c_ms, _ = timed(lambda: w.check_configs(arm[:nc], base[:nc], clearance=True), 3)
        ncl = min(1 << 16, nc)
        o_cl = cchecker.CCloud(cloud).clearance(arm[:ncl].cpu().numpy(), base[:ncl].cpu().numpy(), cap=0.25)
        extras["with_clearance"] = {"value": nc / (c_ms * 1e-3), "unit": UNIT, "configs": nc, "ms": c_ms,
                                    "max_abs_err_vs_fp64_oracle_m": float(np.abs(c_clear[:ncl].cpu().numpy() - o_cl).max()), "cap_m": 0.25}
        del f_arm, f_base, f_valid

    pk = peaks()
    props = torch.cuda.get_device_properties(dev)
    sms = props.multi_processor_count
    fp32_peak = sms * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12      # TFLOP/s, non-tensor
    pairs = max(tim["launch_pairs"], 1)
    k1_ms = tim["phase1_ms"] / pairs                               # avg k_check_configs launch
    k2_ms = tim["phase2_ms"] / pairs                               # avg k_resolve launch
    cfg_per_launch = B * args.steps / pairs
    per_gpu_rate = B / (ms_per_step * 1e-3)
    achieved = per_gpu_rate * F_CFG / 1e12                         # cfg/s x 4,620 over the WHOLE step
    executed = executed_work(args, args.workload) if (world == 1 and not args.no_extras) else None
    roofline = {
        "kernel": "k_check_configs + k_resolve (fused FK + distance-grid classification, exact narrow phase): the whole step",
        "bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
        "definition": "SURVEY.md 8(d): configs/s per GPU x 4,620 algorithmic flop/config / non-tensor FP32 peak; early exit means most of that "
                      "work is never executed - see 'executed' and 'binding_resource'",
        "peak_source": f"{sms} SMs x 128 lanes x 2 flop x {pk['sm_max_mhz']:.0f} MHz ({pk['source']} sm_max_mhz)",
        "algorithmic_flop_per_config": F_CFG, "configs_per_launch": cfg_per_launch,
        "phase1_avg_launch_ms": k1_ms, "phase2_avg_launch_ms": k2_ms,
        "phase1_share_of_step": tim["phase1_ms"] / (ms_per_step * args.steps), "phase2_share_of_step": tim["phase2_ms"] / (ms_per_step * args.steps),
        "traffic": traffic_from_profile(cfg_per_launch, args.vpm),
        "hbm": {"achieved": per_gpu_rate * B_CFG / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": per_gpu_rate * B_CFG / 1e9 / pk["hbm_gbs"],
                "algorithmic_bytes_per_config": B_CFG, "peak_source": pk["source"]},
        "executed": executed,
    }
    if executed and "probes_per_config" in executed:
        probes_s = executed["probes_per_config"] * cfg_per_launch / (k1_ms * 1e-3)
        req_peak = sms * pk["sm_max_mhz"] * 1e6
        roofline["binding_resource"] = {
            "name": "phase 1: stage A is bound by dependent probe latencies, stage B by instruction issue (ncu: smsp__issue_active 77 %, "
                    "18.6 of 32 lanes active; DESIGN.md 5.2 / 8); the L1-miss request rate (one 32-byte sector request per SM per clock, every "
                    "probe of a lane its own sector) that bound round 1's kernel at 0.65 is the fraction reported here",
            "achieved_probes_per_s": probes_s, "peak_requests_per_s": req_peak, "frac": probes_s / req_peak,
            "counter": "l1tex__m_l1tex2xbar_req_cycles_active 46 %, smsp__issue_active (profiles/r06z_ncu_full_summary.txt)",
        }
        # executed work: 18 flop sphere transform + 20 flop probe per executed probe, 401 flop of arm-chain frames for the
        # configurations that reach stage B, 9 flop per narrow-phase point test
        ex_flop = 38.0 * executed["probes_per_config"] + 401.0 * executed["stage_b_fraction"] + F_TEST * executed["narrow_point_tests_per_config"]
        roofline["executed_flop_per_config"] = ex_flop
        roofline["executed_flop_frac"] = per_gpu_rate * ex_flop / 1e12 / fp32_peak
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": config, "e2e": e2e, "gpu_launches": cnt["launches"], "clocks": clocks,
        "roofline": roofline, "parity_sample": parity,
        "world": {"build_ms": t_build, "grid": [stats["nx"], stats["ny"], stats["nz"]],
                  "slack_m": stats["slack"], "free_fraction": free_frac, "floor_fraction_free_for_tucked_robot": floor_free,
                  "narrow_items_per_config": cnt["narrow_items"] / (B * args.steps),
                  "definite_collision_fraction": cnt["definite_collisions"] / (B * args.steps),
                  "overflow_inline": cnt["overflow_inline"]},
    }
    line.update(extras)
    if multi:
        line["multi_gpu"] = multi
    if world == 1 and not args.no_cpu:
        r = cpu_arm(args, cloud, 1, 1)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "one_core")}
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
