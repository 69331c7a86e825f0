"""Multi-GPU plumbing: one process per GPU, configurations / edges / robot states sharded by rank.

The path has no data-path collective: every unit (configuration, edge, frame) is independent given
the cloud structure.  The only exchanges are
  1. the world hand-off: the rank that ingested the cloud builds the structure and broadcasts it
     (or broadcasts the raw cloud and every rank rebuilds — ``broadcast_cloud``), and
  2. the optional gather of per-unit verdict bytes back to one rank.
Both go through ``torch.distributed`` (NCCL on GPUs; the same code runs on gloo for the CPU tests,
where ``check`` is any callable — the tests use a stand-in, the product passes ``CollisionWorld``).
"""

from __future__ import annotations

from typing import Callable

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block split: rank r owns [r*total/world, (r+1)*total/world) (sizes differ by <= 1)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(int(total), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def broadcast_cloud(points: torch.Tensor | None, src: int = 0, device: torch.device | str = "cpu") -> torch.Tensor:
    """Broadcast a raw (N,3) f32 cloud from ``src`` to every rank (6 MB for 500k points)."""
    rank = dist.get_rank()
    n = torch.tensor([points.shape[0] if rank == src else 0], dtype=torch.int64, device=device)
    dist.broadcast(n, src)
    if rank == src:
        buf = points.to(device=device, dtype=torch.float32).contiguous()
    else:
        buf = torch.empty(int(n.item()), 3, dtype=torch.float32, device=device)
    if buf.numel():
        dist.broadcast(buf, src)
    return buf


def broadcast_world(world_obj, src: int = 0) -> None:
    """Broadcast the BUILT structure of a ``CollisionWorld`` from ``src`` into every rank's env."""
    rank = dist.get_rank()
    dev = world_obj.device
    n = torch.tensor([world_obj.used_bytes() if rank == src else 0], dtype=torch.int64, device=dev)
    dist.broadcast(n, src)
    view = world_obj.blob_view(int(n.item()))
    dist.broadcast(view, src)
    if rank != src:
        world_obj.adopt_blob(int(n.item()))


def sharded_verdicts(check: Callable[[torch.Tensor, torch.Tensor], torch.Tensor], arm: torch.Tensor, base: torch.Tensor,
                     gather: bool = True) -> torch.Tensor:
    """Each rank checks its block of the SAME global (B,8)/(B,3) batch; verdict bytes are all-gathered.

    Returns the full (B,) uint8 verdict vector on every rank (``gather=True``) or the local block.
    """
    rank, world = dist.get_rank(), dist.get_world_size()
    B = arm.shape[0]
    lo, hi = shard_range(B, rank, world)
    local = check(arm[lo:hi], base[lo:hi]).to(torch.uint8)
    if not gather:
        return local
    width = (B + world - 1) // world                      # equal-size slots for all_gather
    slot = torch.zeros(width, dtype=torch.uint8, device=local.device)
    slot[: hi - lo] = local
    out = [torch.empty_like(slot) for _ in range(world)]
    dist.all_gather(out, slot)
    parts = []
    for r in range(world):
        a, b = shard_range(B, r, world)
        parts.append(out[r][: b - a])
    return torch.cat(parts)
