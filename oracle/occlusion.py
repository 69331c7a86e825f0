"""Self-occlusion check: do the projected robot spheres overlap the projected target sphere?

Oracle (test infrastructure).  Follows ``TyGrit/checker/occlusion.py:24-116`` line by line, with
``cv2.circle(mask, (u, v), r, 1, -1)`` replaced by a restatement of OpenCV's filled midpoint circle
(``modules/imgproc/src/drawing.cpp``, ``Circle``; OpenCV 4.13 is what the reference runs on here) so
that the oracle needs no cv2 on the GPU box.  ``tests/test_oracle.py`` checks the rasteriser against
cv2 itself when cv2 is importable, and the whole function against vectors produced by the unmodified
reference (``tools/make_golden_occlusion.py`` -> ``tests/golden/occlusion.npz``): PINNED.
"""

from __future__ import annotations

import numpy as np

from . import fetch_model as fm
from . import fk as ofk

# occlusion.py:21  _R_LINK_TO_CV = inv(R_CV_TO_CAMERA_LINK): cv.x = -link.y, cv.y = -link.z, cv.z = link.x
R_LINK_TO_CV = np.array([[0.0, -1.0, 0.0], [0.0, 0.0, -1.0], [1.0, 0.0, 0.0]])


def disc_half_widths(radius: int) -> np.ndarray:
    """Half-width of the filled circle in the row ``d`` pixels from the centre, ``d = 0..radius``.

    OpenCV's ``Circle``: starting from ``(dx, dy) = (radius, 0)`` every step fills rows ``cy -+ dy`` over
    ``cx -+ dx`` and rows ``cy -+ dx`` over ``cx -+ dy``, then ``dy += 1; err += plus; plus += 2`` and, when
    ``err > 0``, ``err -= minus; dx -= 1; minus -= 2``.
    """
    hw = np.full(radius + 1, -1, dtype=np.int64)
    err, dx, dy, plus, minus = 0, radius, 0, 1, 2 * radius - 1
    while dx >= dy:
        hw[dy] = max(hw[dy], dx)
        hw[dx] = max(hw[dx], dy)
        dy += 1
        err += plus
        plus += 2
        if err > 0:
            err -= minus
            dx -= 1
            minus -= 2
    return hw


def draw_disc(mask: np.ndarray, u: int, v: int, radius: int) -> None:
    """``cv2.circle(mask, (u, v), radius, 1, -1)`` clipped to the image."""
    h, w = mask.shape
    hw = disc_half_widths(radius)
    for d in range(radius + 1):
        if hw[d] < 0:
            continue
        x0, x1 = max(u - hw[d], 0), min(u + hw[d], w - 1)
        if x0 > x1:
            continue
        for y in (v - d, v + d):
            if 0 <= y < h:
                mask[y, x0:x1 + 1] = 1


def spheres_to_mask(centers_cam, radii, K, h: int, w: int) -> np.ndarray:
    """occlusion.py:24-45."""
    mask = np.zeros((h, w), dtype=np.uint8)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    for c, r in zip(centers_cam, radii):
        if c[2] <= 0:
            continue
        u = int(fx * c[0] / c[2] + cx)
        v = int(fy * c[1] / c[2] + cy)
        pr = max(1, int(fx * r / c[2]))
        if 0 <= u < w and 0 <= v < h:
            draw_disc(mask, u, v, pr)
    return mask.astype(bool)


def check_self_occlusion(q10, target_center, target_radius: float, K, image_shape) -> bool:
    """occlusion.py:48-116 for the link poses of ``q10`` (fp64 FK of ``oracle.fk``)."""
    R, t = ofk.link_frames(np.asarray(q10, dtype=np.float64)[None])
    R, t = R[0], t[0]
    ht = fm.LINK_NAMES.index("head_tilt_link")
    T = np.eye(4)
    T[:3, :3] = R[ht]
    T[:3, 3] = t[ht] + R[ht] @ np.asarray(fm.HEAD_CAMERA_OFFSET, dtype=np.float64)
    T_cv_base = np.eye(4)
    T_cv_base[:3, :3] = R_LINK_TO_CV
    T_cv_base = T_cv_base @ np.linalg.inv(T)
    Rc, tc = T_cv_base[:3, :3], T_cv_base[:3, 3]
    sph = np.asarray(fm.SPHERE_LOCAL, dtype=np.float64)
    link = np.asarray(fm.SPHERE_LINK)
    centres_base = np.einsum("nij,nj->ni", R[link], sph[:, :3]) + t[link]
    centres_cam = centres_base @ Rc.T + tc
    target_cam = Rc @ np.asarray(target_center, dtype=np.float64) + tc
    h, w = image_shape
    robot = spheres_to_mask(centres_cam, sph[:, 3], K, h, w)
    target = spheres_to_mask(target_cam[None], np.array([target_radius]), K, h, w)
    return bool(np.any(robot & target))
