"""Fetch robot model, by value (oracle copy — test infrastructure).

Numbers follow ``TyGrit/kinematics/fetch/constants.py``:
link offsets ``:17-31``, head camera ``:34-50``, joint limits ``:74-79``,
collision spheres ``:84-225`` (14 links, 111 spheres ``[x, y, z, r]`` in the
link frame).  The kinematic tree (parents / joint axes) follows the chain in
``TyGrit/kinematics/fetch/fk_numpy.py:65-155``.

Kept deliberately independent of ``tygrit_b200/fetch_model.py`` so that a typo
in the product's table is caught by the parity tests.
"""

from __future__ import annotations

import numpy as np

# (link, parent, offset xyz, joint kind, index into q[10])
#   kind: "-" fixed, "pz" prismatic along z, "rx"/"ry"/"rz" revolute
# q = [torso, sh_pan, sh_lift, up_roll, el_flex, fo_roll, wr_flex, wr_roll,
#      head_pan, head_tilt]                      (fk_numpy.py:38-41)
TREE = (
    ("base_link", -1, (0.0, 0.0, 0.0), "-", -1),
    ("torso_fixed_link", 0, (-0.086875, 0.0, 0.377425), "-", -1),
    ("torso_lift_link", 0, (-0.086875, 0.0, 0.37743), "pz", 0),
    ("head_pan_link", 2, (0.053125, 0.0, 0.603001), "rz", 8),
    ("head_tilt_link", 3, (0.14253, 0.0, 0.057999), "ry", 9),
    ("shoulder_pan_link", 2, (0.119525, 0.0, 0.34858), "rz", 1),
    ("shoulder_lift_link", 5, (0.117, 0.0, 0.06), "ry", 2),
    ("upperarm_roll_link", 6, (0.219, 0.0, 0.0), "rx", 3),
    ("elbow_flex_link", 7, (0.133, 0.0, 0.0), "ry", 4),
    ("forearm_roll_link", 8, (0.197, 0.0, 0.0), "rx", 5),
    ("wrist_flex_link", 9, (0.1245, 0.0, 0.0), "ry", 6),
    ("wrist_roll_link", 10, (0.1385, 0.0, 0.0), "rx", 7),
    ("gripper_link", 11, (0.16645, 0.0, 0.0), "-", -1),
    ("r_gripper_finger_link", 12, (0.0, 0.065425, 0.0), "-", -1),
    ("l_gripper_finger_link", 12, (0.0, -0.065425, 0.0), "-", -1),
)
LINK_NAMES = tuple(t[0] for t in TREE)
LINK_INDEX = {n: i for i, n in enumerate(LINK_NAMES)}

HEAD_CAMERA_OFFSET = np.array([0.055, 0.0, 0.0225])
# OpenCV camera axes expressed in head_camera_link axes (constants.py:41-50)
R_CV_TO_CAMERA_LINK = np.array(
    [[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]]
)

JOINT_LOWER = np.array([0.0, -1.6056, -1.221, -np.pi, -2.251, -np.pi, -2.16, -np.pi])
JOINT_UPPER = np.array([0.38615, 1.6056, 1.518, np.pi, 2.251, np.pi, 2.16, np.pi])
# head limits are not in constants.py; the URDF the reference ships has
# pan +-1.57, tilt -0.76..1.45 (resources/fetch/fetch.urdf)
HEAD_LOWER = np.array([-1.57, -0.76])
HEAD_UPPER = np.array([1.57, 1.45])

# Sphere table in the reference's dict order (constants.py:84-225).
# "x,y,z,r;" per sphere.
_SPHERES_TXT = """
base_link: -0.12,0,0.182,0.24; 0.225,0,0.31,0.066; 0.08,-0.06,0.16,0.22;
 0.215,-0.07,0.31,0.066; 0.185,-0.135,0.31,0.066; 0.13,-0.185,0.31,0.066;
 0.065,-0.2,0.31,0.066; 0.01,-0.2,0.31,0.066; 0.08,0.06,0.16,0.22;
 0.215,0.07,0.31,0.066; 0.185,0.135,0.31,0.066; 0.13,0.185,0.31,0.066;
 0.065,0.2,0.31,0.066; 0.01,0.2,0.31,0.066;
torso_lift_link: -0.1,-0.05,0.15,0.15; -0.1,0.05,0.15,0.15; -0.1,0.05,0.3,0.15;
 -0.1,0.05,0.45,0.15; -0.1,-0.05,0.45,0.15; -0.1,-0.05,0.3,0.15; 0.1,0,0.24,0.07;
torso_fixed_link: -0.1,-0.07,0.35,0.12; -0.1,0.07,0.35,0.12; -0.1,-0.07,0.2,0.12;
 -0.1,0.07,0.2,0.12; -0.1,0.07,0.07,0.12; -0.1,-0.07,0.07,0.12;
head_pan_link: 0,0,0.06,0.15; 0.145,0,0.058,0.05; 0.145,-0.0425,0.058,0.05;
 0.145,0.0425,0.058,0.05; 0.145,0.085,0.058,0.05; 0.145,-0.085,0.058,0.05;
 0.0625,-0.115,0.03,0.03; 0.088,-0.115,0.03,0.03; 0.1135,-0.115,0.03,0.03;
 0.139,-0.115,0.03,0.03; 0.0625,-0.115,0.085,0.03; 0.088,-0.115,0.085,0.03;
 0.1135,-0.115,0.085,0.03; 0.139,-0.115,0.085,0.03; 0.16,-0.115,0.075,0.03;
 0.168,-0.115,0.0575,0.03; 0.16,-0.115,0.04,0.03; 0.0625,0.115,0.03,0.03;
 0.088,0.115,0.03,0.03; 0.1135,0.115,0.03,0.03; 0.139,0.115,0.03,0.03;
 0.0625,0.115,0.085,0.03; 0.088,0.115,0.085,0.03; 0.1135,0.115,0.085,0.03;
 0.139,0.115,0.085,0.03; 0.16,0.115,0.075,0.03; 0.168,0.115,0.0575,0.03;
 0.16,0.115,0.04,0.03;
shoulder_pan_link: 0,0,0,0.055; 0.025,-0.015,0.035,0.055; 0.05,-0.03,0.06,0.055;
 0.12,-0.03,0.06,0.055;
shoulder_lift_link: 0.025,0.04,0.025,0.04; -0.025,0.04,-0.025,0.04;
 0.025,0.04,-0.025,0.04; -0.025,0.04,0.025,0.04; 0.08,0,0,0.055; 0.11,0,0,0.055;
 0.14,0,0,0.055;
upperarm_roll_link: -0.02,0,0,0.055; 0.03,0,0,0.055; 0.08,0,0,0.055;
 0.11,-0.045,0.02,0.03; 0.11,-0.045,-0.02,0.03; 0.155,-0.045,0.02,0.03;
 0.155,-0.045,-0.02,0.03; 0.13,0,0,0.055;
elbow_flex_link: 0.02,0.045,0.02,0.03; 0.02,0.045,-0.02,0.03; -0.02,0.045,0.02,0.03;
 -0.02,0.045,-0.02,0.03; 0.08,0,0,0.055; 0.14,0,0,0.055;
forearm_roll_link: 0,0,0,0.055; 0.05,-0.06,0.02,0.03; 0.05,-0.06,-0.02,0.03;
 0.1,-0.06,0.02,0.03; 0.1,-0.06,-0.02,0.03; 0.15,-0.06,0.02,0.03;
 0.15,-0.06,-0.02,0.03;
wrist_flex_link: 0,0,0,0.055; 0.06,0,0,0.055; 0.02,0.045,0.02,0.03;
 0.02,0.045,-0.02,0.03; -0.02,0.045,0.02,0.03; -0.02,0.045,-0.02,0.03;
wrist_roll_link: -0.03,0,0,0.055; 0,0,0,0.055;
gripper_link: -0.07,0.02,0,0.05; -0.07,-0.02,0,0.05; -0.1,0.02,0,0.05;
 -0.1,-0.02,0,0.05;
r_gripper_finger_link: 0.017,-0.0085,-0.005,0.012; 0.017,-0.0085,0.005,0.012;
 0,-0.0085,-0.005,0.012; 0,-0.0085,0.005,0.012; -0.017,-0.0085,-0.005,0.012;
 -0.017,-0.0085,0.005,0.012;
l_gripper_finger_link: 0.017,0.0085,-0.005,0.012; 0.017,0.0085,0.005,0.012;
 0,0.0085,-0.005,0.012; 0,0.0085,0.005,0.012; -0.017,0.0085,-0.005,0.012;
 -0.017,0.0085,0.005,0.012;
"""


def _parse_spheres(txt: str) -> dict[str, np.ndarray]:
    out: dict[str, np.ndarray] = {}
    name = None
    rows: list[list[float]] = []
    for tok in txt.replace("\n", " ").split(";"):
        tok = tok.strip()
        if not tok:
            continue
        if ":" in tok:
            if name is not None:
                out[name] = np.array(rows, dtype=np.float64)
            name, tok = (s.strip() for s in tok.split(":"))
            rows = []
        rows.append([float(v) for v in tok.split(",")])
    out[name] = np.array(rows, dtype=np.float64)
    return out


SPHERES_BY_LINK = _parse_spheres(_SPHERES_TXT)          # dict order = reference order
SPHERE_LINK = np.concatenate(
    [np.full(len(v), LINK_INDEX[k]) for k, v in SPHERES_BY_LINK.items()]
).astype(np.int64)                                       # (111,) link index per sphere
SPHERE_LOCAL = np.vstack(list(SPHERES_BY_LINK.values()))  # (111, 4) x,y,z,r
N_SPHERES = SPHERE_LOCAL.shape[0]
assert N_SPHERES == 111

# Non-adjacent link pairs that can physically self-collide (TyGrit/rl/train.py:78-107,
# SELF_COLLISION_PAIRS).  head_tilt_link carries no spheres, so its eight pairs contribute
# no sphere-sphere test; the remaining 18 link pairs expand to 1,764 sphere pairs.
SELF_COLLISION_PAIRS = tuple(
    [("head_pan_link", b) for b in ("upperarm_roll_link", "elbow_flex_link", "forearm_roll_link", "wrist_flex_link",
                                    "wrist_roll_link", "gripper_link", "l_gripper_finger_link", "r_gripper_finger_link")]
    + [("head_tilt_link", b) for b in ("upperarm_roll_link", "elbow_flex_link", "forearm_roll_link", "wrist_flex_link",
                                       "wrist_roll_link", "gripper_link", "l_gripper_finger_link", "r_gripper_finger_link")]
    + [(a, b) for a in ("base_link", "torso_lift_link")
       for b in ("wrist_flex_link", "wrist_roll_link", "gripper_link", "l_gripper_finger_link", "r_gripper_finger_link")]
)


def _self_sphere_pairs():
    ii, jj = [], []
    for a, b in SELF_COLLISION_PAIRS:
        ia = np.flatnonzero(SPHERE_LINK == LINK_INDEX[a])
        ib = np.flatnonzero(SPHERE_LINK == LINK_INDEX[b])
        for i in ia:
            for j in ib:
                ii.append(i)
                jj.append(j)
    return np.asarray(ii, dtype=np.int64), np.asarray(jj, dtype=np.int64)


SELF_PAIR_I, SELF_PAIR_J = _self_sphere_pairs()        # (1764,) sphere indices (reference order)
assert SELF_PAIR_I.shape[0] == 1764
