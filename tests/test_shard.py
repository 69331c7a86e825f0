"""Multi-rank host logic on CPU: world_size-2 gloo process group.

The product passes ``CollisionWorld.check_configs`` as the ``check`` callable; here a deterministic
stand-in (a pure function of the configuration) plays that role so that the sharding, the ragged
block split and the verdict gather can be verified without a GPU.
"""

from __future__ import annotations

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tygrit_b200 import shard


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _stand_in_check(arm: torch.Tensor, base: torch.Tensor) -> torch.Tensor:
    return ((arm.sum(dim=1) + base[:, 0]) > 0).to(torch.uint8)


def _worker(rank: int, world: int, port: int, B: int, out_dir: str) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        arm = torch.randn(B, 8, generator=g)
        base = torch.randn(B, 3, generator=g)
        full = shard.sharded_verdicts(_stand_in_check, arm, base, gather=True)
        local = shard.sharded_verdicts(_stand_in_check, arm, base, gather=False)
        lo, hi = shard.shard_range(B, rank, world)
        assert local.shape[0] == hi - lo
        assert torch.equal(full, _stand_in_check(arm, base))
        assert torch.equal(full[lo:hi], local)
        # raw-cloud broadcast: rank 0 owns the cloud, the others receive it
        cloud = torch.arange(30, dtype=torch.float32).reshape(10, 3) if rank == 0 else None
        got = shard.broadcast_cloud(cloud, src=0)
        assert got.shape == (10, 3) and torch.equal(got, torch.arange(30, dtype=torch.float32).reshape(10, 3))
        empty = shard.broadcast_cloud(torch.empty(0, 3) if rank == 0 else None, src=0)
        assert empty.shape == (0, 3)
        np.save(os.path.join(out_dir, f"ok{rank}.npy"), np.array([1]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [1001, 2, 1])
def test_sharded_verdicts_gloo_world2(tmp_path, B):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), B, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}.npy") for r in range(world))


def test_shard_range_covers_everything():
    for total in (0, 1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            blocks = [shard.shard_range(total, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(blocks[:-1], blocks[1:]))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(10, 2, 2)
