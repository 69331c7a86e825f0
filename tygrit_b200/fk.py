"""Batched Fetch FK solver — drop-in for the reference's ``"fk_batch"`` solver.

``B200BatchFK`` satisfies the ``BatchFKSolver`` protocol (``TyGrit/kinematics/fk.py:84-106``):
``solve((B,10) tensor) -> {link: (B,4,4) tensor}`` in ``base_link``, same 15 link names as
``TyGrit/kinematics/fetch/fk_torch.py:119-283``, computed by one CUDA kernel instead of ~25 bmm.
"""

from __future__ import annotations

import torch

from . import engine
from .fetch_model import LINK_NAMES


class B200BatchFK:
    @property
    def base_frame(self) -> str:
        return "base_link"

    def solve(self, joint_angles: torch.Tensor) -> dict[str, torch.Tensor]:
        if joint_angles.ndim != 2 or joint_angles.shape[1] != 10:
            raise ValueError(f"expected (B, 10) joint tensor, got {tuple(joint_angles.shape)}")
        if not joint_angles.is_cuda:
            raise RuntimeError("B200BatchFK needs a CUDA tensor (no CPU fallback)")
        frames = engine.fk_frames(joint_angles)
        return {name: frames[:, i] for i, name in enumerate(LINK_NAMES)}


def batch_forward_kinematics(joint_angles: torch.Tensor) -> dict[str, torch.Tensor]:
    """Function form, same name as ``fk_torch.batch_forward_kinematics``."""
    return B200BatchFK().solve(joint_angles)


def create_fk_solver(robot: str, solver: str) -> B200BatchFK:
    """Factory branch for ``TyGrit/kinematics/fk.py:109-133`` / ``fetch/fk.py:25-50``."""
    if robot != "fetch":
        raise ValueError(f"Unknown robot: {robot}")
    if solver in ("fk_b200", "fk_batch"):
        return B200BatchFK()
    raise ValueError(f"Unknown Fetch FK solver: {solver!r}. Available: fk_b200")
