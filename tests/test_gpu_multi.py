"""GPU, >= 2 devices: the multi-rank path on NCCL with the REAL checker — rank 0 builds the structure,
``shard.broadcast_world`` moves the blob over NCCL into every rank's env, every rank checks its block of the same global
batch with ``CollisionWorld.check_configs`` and ``shard.sharded_verdicts`` all-gathers the verdict bytes; the gathered
vector must equal the fp64 oracle's verdicts (outside the band) on every rank.  Skipped on a one-GPU box (the driver's
GPU test tier); run with ``gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`` (log under profiles/).
"""

from __future__ import annotations

import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank: int, world: int, port: int, out_dir: str) -> None:
    import torch.distributed as dist

    from oracle import cchecker
    from tygrit_b200 import shard, synth
    from tygrit_b200.engine import CollisionWorld

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        cloud = synth.room_cloud(120_000)
        w = CollisionWorld(device=rank, max_points=1 << 18)
        if rank == 0:
            w.build(torch.from_numpy(cloud).to(dev))
        shard.broadcast_world(w, src=0)                              # NCCL broadcast of the built structure
        arm, base = synth.random_configs(300_001, seed=9)           # ragged split; >= 2^18 per call exercises the coarse variant on rank 0
        a, b = torch.from_numpy(arm).to(dev), torch.from_numpy(base).to(dev)
        full = shard.sharded_verdicts(lambda x, y: w.check_configs(x, y), a, b, gather=True)
        assert full.shape == (arm.shape[0],) and full.dtype == torch.uint8
        cl = cchecker.CCloud(cloud).clearance(arm[:60_000], base[:60_000], cap=1e-3)
        d = np.abs(cl) > 1e-5
        got = full[:60_000].cpu().numpy().astype(bool)
        assert np.array_equal(got[d], (cl > 0)[d]), f"rank {rank}: gathered verdicts differ from the oracle"
        # the tail block (owned by the last rank) against a single-GPU check on this rank
        lo, hi = shard.shard_range(arm.shape[0], world - 1, world)
        own = w.check_configs(a[lo:hi], b[lo:hi])
        assert torch.equal(full[lo:hi], own)
        # raw-cloud broadcast + per-rank rebuild gives the same world
        c2 = shard.broadcast_cloud(torch.from_numpy(cloud) if rank == 0 else None, src=0, device=dev)
        w2 = CollisionWorld(device=rank, max_points=1 << 18)
        w2.build(c2)
        assert torch.equal(w2.check_configs(a[:50_000], b[:50_000]), full[:50_000])
        np.save(os.path.join(out_dir, f"ok{rank}.npy"), np.array([1]))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL refuses two ranks on one device)")
def test_sharded_verdicts_nccl_world2(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}.npy") for r in range(2))
