"""Fetch collision model used by the B200 engine (product side).

Single source of truth for the numbers the CUDA kernels bake in:
``tools/gen_fetch_header.py`` turns this module into
``tygrit_b200/csrc/fetch_model.cuh``.  Values are the robot's URDF data as
listed in the reference (``TyGrit/kinematics/fetch/constants.py:17-31`` link
offsets, ``:74-79`` joint limits, ``:84-225`` collision spheres).

Layout notes
------------
* ``LINKS``: kinematic tree in evaluation order (parents first).
* ``SPHERES``: (111, 5) rows ``[link_index, x, y, z, r]`` in the reference's
  sphere order (its ``FETCH_SPHERES`` dict order) — the order of
  ``sphere_centres`` outputs and of the self-filter centre list.
* ``link_bounds()``: one enclosing sphere per link for the coarse pass.
"""

from __future__ import annotations

import numpy as np

FIXED, PRISM_Z, REV_X, REV_Y, REV_Z = 0, 1, 2, 3, 4

# name, parent, (ox, oy, oz), joint kind, joint slot in q[10]
LINKS = [
    ("base_link",             -1, (0.0, 0.0, 0.0),             FIXED,   -1),
    ("torso_fixed_link",       0, (-0.086875, 0.0, 0.377425),  FIXED,   -1),
    ("torso_lift_link",        0, (-0.086875, 0.0, 0.37743),   PRISM_Z,  0),
    ("head_pan_link",          2, (0.053125, 0.0, 0.603001),   REV_Z,    8),
    ("head_tilt_link",         3, (0.14253, 0.0, 0.057999),    REV_Y,    9),
    ("shoulder_pan_link",      2, (0.119525, 0.0, 0.34858),    REV_Z,    1),
    ("shoulder_lift_link",     5, (0.117, 0.0, 0.06),          REV_Y,    2),
    ("upperarm_roll_link",     6, (0.219, 0.0, 0.0),           REV_X,    3),
    ("elbow_flex_link",        7, (0.133, 0.0, 0.0),           REV_Y,    4),
    ("forearm_roll_link",      8, (0.197, 0.0, 0.0),           REV_X,    5),
    ("wrist_flex_link",        9, (0.1245, 0.0, 0.0),          REV_Y,    6),
    ("wrist_roll_link",       10, (0.1385, 0.0, 0.0),          REV_X,    7),
    ("gripper_link",          11, (0.16645, 0.0, 0.0),         FIXED,   -1),
    ("r_gripper_finger_link", 12, (0.0, 0.065425, 0.0),        FIXED,   -1),
    ("l_gripper_finger_link", 12, (0.0, -0.065425, 0.0),       FIXED,   -1),
]
LINK_NAMES = tuple(l[0] for l in LINKS)
N_LINKS = len(LINKS)
GRIPPER_LINK = LINK_NAMES.index("gripper_link")
HEAD_TILT_LINK = LINK_NAMES.index("head_tilt_link")

HEAD_CAMERA_OFFSET = (0.055, 0.0, 0.0225)

PLANNING_JOINT_NAMES = (
    "torso_lift_joint", "shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint",
    "elbow_flex_joint", "forearm_roll_joint", "wrist_flex_joint", "wrist_roll_joint",
)
JOINT_LOWER = np.array([0.0, -1.6056, -1.221, -np.pi, -2.251, -np.pi, -2.16, -np.pi])
JOINT_UPPER = np.array([0.38615, 1.6056, 1.518, np.pi, 2.251, np.pi, 2.16, np.pi])
HEAD_LOWER = np.array([-1.57, -0.76])
HEAD_UPPER = np.array([1.57, 1.45])


def _rows(link: str, xyzr) -> list[list[float]]:
    li = LINK_NAMES.index(link)
    return [[li, *row] for row in xyzr]


def _mirror_y(rows):
    return [[x, -y, z, r] for x, y, z, r in rows]


# Collision spheres, reference order.  Symmetric groups are written once and
# mirrored so the table stays reviewable.
_BASE_RING = [[0.215, -0.07, 0.31, 0.066], [0.185, -0.135, 0.31, 0.066], [0.13, -0.185, 0.31, 0.066],
              [0.065, -0.2, 0.31, 0.066], [0.01, -0.2, 0.31, 0.066]]
_HEAD_EAR = [[0.0625, -0.115, 0.03, 0.03], [0.088, -0.115, 0.03, 0.03], [0.1135, -0.115, 0.03, 0.03],
             [0.139, -0.115, 0.03, 0.03], [0.0625, -0.115, 0.085, 0.03], [0.088, -0.115, 0.085, 0.03],
             [0.1135, -0.115, 0.085, 0.03], [0.139, -0.115, 0.085, 0.03], [0.16, -0.115, 0.075, 0.03],
             [0.168, -0.115, 0.0575, 0.03], [0.16, -0.115, 0.04, 0.03]]
_FINGER = [[0.017, -0.0085, -0.005, 0.012], [0.017, -0.0085, 0.005, 0.012], [0.0, -0.0085, -0.005, 0.012],
           [0.0, -0.0085, 0.005, 0.012], [-0.017, -0.0085, -0.005, 0.012], [-0.017, -0.0085, 0.005, 0.012]]

_S: list[list[float]] = []
_S += _rows("base_link", [[-0.12, 0.0, 0.182, 0.24], [0.225, 0.0, 0.31, 0.066], [0.08, -0.06, 0.16, 0.22]]
            + _BASE_RING + [[0.08, 0.06, 0.16, 0.22]] + _mirror_y(_BASE_RING))
_S += _rows("torso_lift_link", [[-0.1, -0.05, 0.15, 0.15], [-0.1, 0.05, 0.15, 0.15], [-0.1, 0.05, 0.3, 0.15],
                                [-0.1, 0.05, 0.45, 0.15], [-0.1, -0.05, 0.45, 0.15], [-0.1, -0.05, 0.3, 0.15],
                                [0.1, 0.0, 0.24, 0.07]])
_S += _rows("torso_fixed_link", [[-0.1, -0.07, 0.35, 0.12], [-0.1, 0.07, 0.35, 0.12], [-0.1, -0.07, 0.2, 0.12],
                                 [-0.1, 0.07, 0.2, 0.12], [-0.1, 0.07, 0.07, 0.12], [-0.1, -0.07, 0.07, 0.12]])
_S += _rows("head_pan_link", [[0.0, 0.0, 0.06, 0.15], [0.145, 0.0, 0.058, 0.05], [0.145, -0.0425, 0.058, 0.05],
                              [0.145, 0.0425, 0.058, 0.05], [0.145, 0.085, 0.058, 0.05],
                              [0.145, -0.085, 0.058, 0.05]] + _HEAD_EAR + _mirror_y(_HEAD_EAR))
_S += _rows("shoulder_pan_link", [[0.0, 0.0, 0.0, 0.055], [0.025, -0.015, 0.035, 0.055],
                                  [0.05, -0.03, 0.06, 0.055], [0.12, -0.03, 0.06, 0.055]])
_S += _rows("shoulder_lift_link", [[0.025, 0.04, 0.025, 0.04], [-0.025, 0.04, -0.025, 0.04],
                                   [0.025, 0.04, -0.025, 0.04], [-0.025, 0.04, 0.025, 0.04],
                                   [0.08, 0.0, 0.0, 0.055], [0.11, 0.0, 0.0, 0.055], [0.14, 0.0, 0.0, 0.055]])
_S += _rows("upperarm_roll_link", [[-0.02, 0.0, 0.0, 0.055], [0.03, 0.0, 0.0, 0.055], [0.08, 0.0, 0.0, 0.055],
                                   [0.11, -0.045, 0.02, 0.03], [0.11, -0.045, -0.02, 0.03],
                                   [0.155, -0.045, 0.02, 0.03], [0.155, -0.045, -0.02, 0.03],
                                   [0.13, 0.0, 0.0, 0.055]])
_S += _rows("elbow_flex_link", [[0.02, 0.045, 0.02, 0.03], [0.02, 0.045, -0.02, 0.03], [-0.02, 0.045, 0.02, 0.03],
                                [-0.02, 0.045, -0.02, 0.03], [0.08, 0.0, 0.0, 0.055], [0.14, 0.0, 0.0, 0.055]])
_S += _rows("forearm_roll_link", [[0.0, 0.0, 0.0, 0.055], [0.05, -0.06, 0.02, 0.03], [0.05, -0.06, -0.02, 0.03],
                                  [0.1, -0.06, 0.02, 0.03], [0.1, -0.06, -0.02, 0.03], [0.15, -0.06, 0.02, 0.03],
                                  [0.15, -0.06, -0.02, 0.03]])
_S += _rows("wrist_flex_link", [[0.0, 0.0, 0.0, 0.055], [0.06, 0.0, 0.0, 0.055], [0.02, 0.045, 0.02, 0.03],
                                [0.02, 0.045, -0.02, 0.03], [-0.02, 0.045, 0.02, 0.03], [-0.02, 0.045, -0.02, 0.03]])
_S += _rows("wrist_roll_link", [[-0.03, 0.0, 0.0, 0.055], [0.0, 0.0, 0.0, 0.055]])
_S += _rows("gripper_link", [[-0.07, 0.02, 0.0, 0.05], [-0.07, -0.02, 0.0, 0.05], [-0.1, 0.02, 0.0, 0.05],
                             [-0.1, -0.02, 0.0, 0.05]])
_S += _rows("r_gripper_finger_link", _FINGER)
_S += _rows("l_gripper_finger_link", _mirror_y(_FINGER))

SPHERES = np.array(_S, dtype=np.float64)          # (111, 5): link, x, y, z, r
N_SPHERES = SPHERES.shape[0]
assert N_SPHERES == 111, N_SPHERES
SPHERE_LINK = SPHERES[:, 0].astype(np.int64)
SPHERE_XYZR = SPHERES[:, 1:5]


# Non-adjacent link pairs that can physically self-collide (reference list: TyGrit/rl/train.py:78-107,
# SELF_COLLISION_PAIRS).  head_tilt_link has no spheres, so its pairs add no sphere-sphere test.
_DISTAL = ("wrist_flex_link", "wrist_roll_link", "gripper_link", "l_gripper_finger_link", "r_gripper_finger_link")
_ARM_FROM_UPPERARM = ("upperarm_roll_link", "elbow_flex_link", "forearm_roll_link") + _DISTAL
SELF_COLLISION_PAIRS = tuple(
    [("head_pan_link", b) for b in _ARM_FROM_UPPERARM] + [("head_tilt_link", b) for b in _ARM_FROM_UPPERARM]
    + [("base_link", b) for b in _DISTAL] + [("torso_lift_link", b) for b in _DISTAL]
)


def fetch_spheres_dict() -> dict[str, list[list[float]]]:
    """The table in the reference's ``FETCH_SPHERES`` shape (link -> [[x,y,z,r], ...])."""
    out: dict[str, list[list[float]]] = {}
    for li, x, y, z, r in SPHERES:
        out.setdefault(LINK_NAMES[int(li)], []).append([float(x), float(y), float(z), float(r)])
    return out


def enclosing_sphere(xyzr: np.ndarray, iters: int = 2000) -> np.ndarray:
    """Small enclosing sphere ``[x,y,z,R]`` of a set of spheres (Badoiu-Clarkson style).

    Any sphere that contains all members is valid for the coarse pass; this one
    is within a fraction of a millimetre of the optimum, which is all that
    matters for culling efficiency.
    """
    c = xyzr[:, :3].mean(axis=0)
    for k in range(1, iters + 1):
        d = np.linalg.norm(xyzr[:, :3] - c, axis=1) + xyzr[:, 3]
        far = int(np.argmax(d))
        dirv = xyzr[far, :3] - c
        n = np.linalg.norm(dirv)
        if n < 1e-12:
            break
        # move towards the farthest member by a shrinking step
        c = c + dirv / n * (d[far] - np.median(d)) / (k + 1)
    R = float((np.linalg.norm(xyzr[:, :3] - c, axis=1) + xyzr[:, 3]).max())
    return np.array([c[0], c[1], c[2], R + 1e-9])


def link_bounds() -> np.ndarray:
    """(N_LINKS, 4) enclosing sphere per link in the link frame (R = 0 if no spheres)."""
    out = np.zeros((N_LINKS, 4))
    for li in range(N_LINKS):
        m = SPHERE_XYZR[SPHERE_LINK == li]
        if len(m):
            out[li] = enclosing_sphere(m)
    return out
