"""fp64 forward kinematics of the Fetch link tree and its 111 collision spheres.

Oracle (test infrastructure).  Follows:
* link chain, joint axes, offsets: ``TyGrit/kinematics/fetch/fk_numpy.py:65-155``
* base pose composition ``Rz(theta)``, ``t=(x, y, 0)``:
  ``TyGrit/utils/transforms.py:116-127``
* sphere centre = ``R_link @ c_local + t_link``:
  ``TyGrit/kinematics/robot_filter.py:35-44``
* camera pose: ``TyGrit/kinematics/fetch/camera.py:25-58``
"""

from __future__ import annotations

import numpy as np

from . import fetch_model as fm


def _axis_rotation(kind: str, a: np.ndarray) -> np.ndarray:
    """(B,) angles -> (B,3,3) rotation about x / y / z (right-handed)."""
    c, s = np.cos(a), np.sin(a)
    z, o = np.zeros_like(a), np.ones_like(a)
    if kind == "rx":
        rows = [[o, z, z], [z, c, -s], [z, s, c]]
    elif kind == "ry":
        rows = [[c, z, s], [z, o, z], [-s, z, c]]
    elif kind == "rz":
        rows = [[c, -s, z], [s, c, z], [z, z, o]]
    else:  # pragma: no cover
        raise ValueError(kind)
    return np.stack([np.stack(r, axis=-1) for r in rows], axis=-2)


def link_frames(q10: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """Link frames in base_link for ``q10`` of shape (B, 10) (fp64).

    Returns ``(R, t)`` with shapes (B, 15, 3, 3) and (B, 15, 3); link order is
    ``fetch_model.LINK_NAMES``.
    """
    q = np.atleast_2d(np.asarray(q10, dtype=np.float64))
    B = q.shape[0]
    R = np.zeros((B, len(fm.TREE), 3, 3))
    t = np.zeros((B, len(fm.TREE), 3))
    eye = np.broadcast_to(np.eye(3), (B, 3, 3))
    for i, (_, parent, off, kind, qi) in enumerate(fm.TREE):
        if parent < 0:
            R[:, i], t[:, i] = eye, 0.0
            continue
        off_v = np.broadcast_to(np.asarray(off, dtype=np.float64), (B, 3)).copy()
        if kind == "pz":
            off_v[:, 2] += q[:, qi]
        t[:, i] = np.einsum("bij,bj->bi", R[:, parent], off_v) + t[:, parent]
        if kind in ("rx", "ry", "rz"):
            R[:, i] = R[:, parent] @ _axis_rotation(kind, q[:, qi])
        else:
            R[:, i] = R[:, parent]
    return R, t


def link_frames_4x4(q10: np.ndarray) -> np.ndarray:
    """(B, 15, 4, 4) homogeneous link frames in base_link."""
    R, t = link_frames(q10)
    T = np.zeros(R.shape[:2] + (4, 4))
    T[..., :3, :3] = R
    T[..., :3, 3] = t
    T[..., 3, 3] = 1.0
    return T


def base_transform(base_xyt: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """SE(2) base pose ``[x, y, theta]`` (B,3) -> (Rz (B,3,3), t (B,3))."""
    b = np.atleast_2d(np.asarray(base_xyt, dtype=np.float64))
    Rz = _axis_rotation("rz", b[:, 2])
    t = np.stack([b[:, 0], b[:, 1], np.zeros(len(b))], axis=-1)
    return Rz, t


def world_link_frames(q10: np.ndarray, base_xyt: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    R, t = link_frames(q10)
    Rb, tb = base_transform(base_xyt)
    Rw = np.einsum("bij,bljk->blik", Rb, R)
    tw = np.einsum("bij,blj->bli", Rb, t) + tb[:, None, :]
    return Rw, tw


def sphere_centres(arm8: np.ndarray, base_xyt: np.ndarray, head2: np.ndarray | None = None) -> np.ndarray:
    """World centres of the 111 collision spheres, (B, 111, 3) fp64.

    ``head2`` defaults to (0, 0): the planner's robot model has the head joints
    fixed (oracle assumption 1).  Sphere order = the reference's
    ``FETCH_SPHERES`` dict order.
    """
    arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
    B = arm.shape[0]
    head = np.zeros((B, 2)) if head2 is None else np.atleast_2d(np.asarray(head2, dtype=np.float64))
    q10 = np.concatenate([arm, head], axis=1)
    Rw, tw = world_link_frames(q10, base_xyt)
    Rs = Rw[:, fm.SPHERE_LINK]                      # (B,111,3,3)
    ts = tw[:, fm.SPHERE_LINK]                      # (B,111,3)
    return np.einsum("bsij,sj->bsi", Rs, fm.SPHERE_LOCAL[:, :3]) + ts


def camera_pose(arm8, head2, base_xyt) -> np.ndarray:
    """cam2world (OpenCV convention), (B,4,4) fp64 — camera.py:25-58."""
    arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
    head = np.atleast_2d(np.asarray(head2, dtype=np.float64))
    Rw, tw = world_link_frames(np.concatenate([arm, head], axis=1), base_xyt)
    i = fm.LINK_INDEX["head_tilt_link"]
    Rt, tt = Rw[:, i], tw[:, i]
    T = np.zeros((arm.shape[0], 4, 4))
    T[:, :3, :3] = Rt @ fm.R_CV_TO_CAMERA_LINK
    T[:, :3, 3] = np.einsum("bij,j->bi", Rt, fm.HEAD_CAMERA_OFFSET) + tt
    T[:, 3, 3] = 1.0
    return T
