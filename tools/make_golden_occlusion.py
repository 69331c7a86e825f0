#!/usr/bin/env python
"""Generate ``tests/golden/occlusion.npz`` by running the REFERENCE's own ``check_self_occlusion`` here.

Build container only (``/root/reference`` and cv2 do not travel): ``python tools/make_golden_occlusion.py``.
Imports the unmodified ``TyGrit.checker.occlusion`` + ``fk_numpy``, evaluates seeded robot states and targets,
stores inputs + verdicts, and asserts on the spot that ``oracle/occlusion.py`` reproduces every verdict and that
its disc rasteriser equals ``cv2.circle`` (the pin of the oracle).
"""

from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("TYGRIT_REFERENCE", "/root/reference"))

import cv2  # noqa: E402
from TyGrit.checker.occlusion import check_self_occlusion  # noqa: E402
from TyGrit.kinematics.fetch.fk_numpy import forward_kinematics  # noqa: E402

from oracle import fetch_model as om  # noqa: E402
from oracle import occlusion as oocc  # noqa: E402

K = np.array([[154.10223, 0.0, 320.0], [0.0, 154.10223, 240.0], [0.0, 0.0, 1.0]])  # test/test_occlusion.py:9-12
SHAPE = (480, 640)


def main() -> None:
    # rasteriser vs cv2.circle: every radius up to 300 at a clipped and an unclipped centre
    for r in list(range(1, 301)) + [517, 1000]:
        for (u, v) in ((320, 240), (3, 470), (639, 0)):
            ref = np.zeros(SHAPE, np.uint8)
            cv2.circle(ref, (u, v), r, 1, -1)
            got = np.zeros(SHAPE, np.uint8)
            oocc.draw_disc(got, u, v, r)
            assert np.array_equal(ref, got), (r, u, v)
    rng = np.random.default_rng(42)
    lo, hi = np.asarray(om.JOINT_LOWER), np.asarray(om.JOINT_UPPER)
    n = 768
    q = np.empty((n, 10))
    q[:, :8] = rng.uniform(lo, hi, (n, 8))
    q[:, 8] = rng.uniform(-1.57, 1.57, n)          # head pan
    q[:, 9] = rng.uniform(-0.76, 1.45, n)          # head tilt
    q[::3, 9] = rng.uniform(0.6, 1.3, len(q[::3]))  # a third of the states look down at the arm
    # targets: in front of the robot, some near the gripper (so that both verdicts occur), some behind / far away
    tgt = np.empty((n, 3))
    tgt[:, 0] = rng.uniform(0.3, 1.6, n)
    tgt[:, 1] = rng.uniform(-0.8, 0.8, n)
    tgt[:, 2] = rng.uniform(0.2, 1.4, n)
    for i in range(0, n, 4):
        g = forward_kinematics(q[i])["gripper_link"]
        tgt[i] = g[:3, 3] + g[:3, :3] @ np.array([0.25, 0.0, 0.0]) + rng.normal(0, 0.05, 3)
    tgt[5::50, 0] = -1.0                           # behind the camera
    rad = rng.uniform(0.02, 0.12, n)
    out = np.zeros(n, bool)
    for i in range(n):
        out[i] = check_self_occlusion(forward_kinematics(q[i]), tgt[i], float(rad[i]), K, SHAPE)
        assert oocc.check_self_occlusion(q[i], tgt[i], float(rad[i]), K, SHAPE) == out[i], i
    # a second, small image (ragged clipping) and a dict without head_tilt_link (-> False, occlusion.py:70-72)
    K2 = K.copy()
    K2[:2] *= 0.25
    out2 = np.array([check_self_occlusion(forward_kinematics(q[i]), tgt[i], float(rad[i]), K2, (120, 160)) for i in range(256)])
    for i in range(256):
        assert oocc.check_self_occlusion(q[i], tgt[i], float(rad[i]), K2, (120, 160)) == out2[i], i
    lp = forward_kinematics(q[0])
    lp.pop("head_tilt_link")
    assert check_self_occlusion(lp, tgt[0], 0.05, K, SHAPE) is False
    path = os.path.join(ROOT, "tests", "golden", "occlusion.npz")
    np.savez_compressed(path, q10=q, target=tgt, radius=rad, K=K, shape=np.array(SHAPE), occluded=out,
                        K_small=K2, shape_small=np.array((120, 160)), occluded_small=out2)
    print("wrote", path, "occluded fraction", out.mean(), out2.mean())


if __name__ == "__main__":
    main()
