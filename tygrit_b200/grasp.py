"""Batched grasp-candidate validation.

The reference's ``GraspSampler.sample`` (``TyGrit/subgoal_generator/samplers/grasp_sampler.py:109-137``) walks
grasps in score order and, per grasp, up to ``max_ik_attempts`` IK solutions, calling ``collision_check(wb)``
once per solution — up to 200 serial Python -> C++ calls — and returns the first solution that passes.  Here
the IK solutions are collected first, ONE ``tg_check_configs`` launch judges all of them, and ``tg_first_valid``
picks the same winner the serial loop would have picked (first grasp, then first attempt, in order).
Grasp prediction and IK stay the caller's (they are injected in the reference too); only the collision part
is this package's business.
"""

from __future__ import annotations

from typing import Callable, Iterable, Sequence

import numpy as np
import torch

from . import _native as N
from .engine import _stream_ptr
from .types import SE2Pose


def first_valid(valid: torch.Tensor, group: int | None = None) -> torch.Tensor:
    """Index of the first non-zero entry of every ``group`` consecutive verdicts ((ceil(B/group),) int32 CUDA,
    -1 where a group has none).  ``group=None``: one group = the whole batch."""
    if not valid.is_cuda:
        raise ValueError("first_valid takes the CUDA verdict tensor of check_configs")
    v = valid.contiguous().view(torch.uint8) if valid.dtype == torch.bool else valid.contiguous()
    if v.dtype != torch.uint8 or v.ndim != 1:
        raise ValueError("valid must be a 1-D uint8/bool tensor")
    B = v.shape[0]
    g = B if group is None else int(group)
    if g < 1:
        if B == 0:
            return torch.empty(0, dtype=torch.int32, device=v.device)
        raise ValueError("group must be >= 1")
    out = torch.empty((B + g - 1) // g, dtype=torch.int32, device=v.device)
    with torch.cuda.device(v.device):
        N.check(N.lib().tg_first_valid(v.data_ptr(), B, g, out.data_ptr(), _stream_ptr(v.device)), "tg_first_valid")
    return out


def validate_grasp_candidates(planner, joints, base_pose: SE2Pose):
    """``joints``: (G, A, 8) array of IK solutions, grasp-major in score order, rows of NaN where IK failed.
    Returns ``(g, a)`` of the first collision-free solution in the reference's visiting order, or ``None``.
    One device launch + one 4-byte read instead of up to G*A ``validate_config`` calls."""
    J = np.asarray(joints, dtype=np.float64)
    if J.ndim != 3 or J.shape[2] != 8:
        raise ValueError("joints must have shape (G, A, 8)")
    G, A = J.shape[:2]
    if G * A == 0:
        return None
    solved = ~np.isnan(J).any(axis=2).reshape(-1)
    arm = np.nan_to_num(J.reshape(-1, 8), nan=0.0).astype(np.float32)
    base = np.tile(np.array([[base_pose.x, base_pose.y, base_pose.theta]], dtype=np.float32), (G * A, 1))
    valid = planner.validate_configs(torch.from_numpy(arm).cuda(), torch.from_numpy(base).cuda())
    valid = valid.view(torch.uint8) if valid.dtype == torch.bool else valid
    valid = valid & torch.from_numpy(solved.astype(np.uint8)).to(valid.device)
    k = int(first_valid(valid).item())
    return None if k < 0 else (k // A, k % A)


class BatchedGraspValidator:
    """IK + collision part of ``GraspSampler.sample`` with the collision calls batched.

    ``ik_solve(T_ee_base, seed) -> (8,) joints`` raises ``ValueError`` when there is no solution, exactly like the
    reference's ``IKSolverBase.solve``.  ``select`` returns what the reference's loop returns: the joints of the
    first (grasp, attempt) that is collision-free, or ``None``.  (The reference stops calling IK at the first
    success; here every candidate's IK runs before the single collision launch.)
    """

    def __init__(self, planner, ik_solve: Callable, max_grasps_to_try: int = 20, max_ik_attempts: int = 10) -> None:
        self._planner, self._ik = planner, ik_solve
        self.max_grasps_to_try, self.max_ik_attempts = int(max_grasps_to_try), int(max_ik_attempts)

    def select(self, ee_poses_base: Sequence[np.ndarray] | Iterable[np.ndarray], base_pose: SE2Pose, seed: np.ndarray):
        poses = list(ee_poses_base)[: self.max_grasps_to_try]
        J = np.full((len(poses), self.max_ik_attempts, 8), np.nan)
        for g, T in enumerate(poses):
            for a in range(self.max_ik_attempts):
                try:
                    J[g, a] = np.asarray(self._ik(T, seed=seed), dtype=np.float64)
                except ValueError:
                    continue
        hit = validate_grasp_candidates(self._planner, J, base_pose)
        return None if hit is None else J[hit[0], hit[1]].copy()
