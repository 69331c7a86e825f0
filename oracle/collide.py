"""Exact sphere-vs-cloud clearance, config verdicts and motion-edge verdicts.

Oracle (test infrastructure).  The arithmetic this restates lives in the
third-party VAMP library (``AdaCompNUS/vamp`` @
``ab0bc393107ebcf445a2e8df52c1aec7835ace4e``), whose source is NOT in the
reference tree — **parity with VAMP is unpinned**.  The contract is therefore
defined here, anchored on TyGrit's call sites:

1. Robot model = 111 spheres of ``FETCH_SPHERES`` with their own radii, posed by
   ``SE2(x, y, theta) . FK(arm[8], head = 0)``
   (``planning/fetch/vamp_preview.py:159-163`` passes ``arm`` + ``[x, y, theta]``).
2. Every cloud point is a sphere of radius ``point_radius`` (default 0.03,
   ``planning/config.py:13``; ``vamp_preview.py:219-224``).
3. ``clearance(config) = min_{s,p} (|c_s - p| - r_s - r_pt)``; collision-free
   iff ``clearance > 0`` (a touching sphere collides).  Configs with
   ``|clearance| <= 1e-5`` m are exempt from bit-exact comparison (north_star).
4. Extra sphere obstacles (``vamp_preview.py:242-255``) and end-effector
   attachments (``:257-304``; spheres given in the attachment frame, which is the
   gripper_link frame composed with ``offset_pos`` / ``offset_quat [x,y,z,w]``)
   take part identically; attachment spheres are also tested against the cloud.
5. No joint-limit term.  The self-collision term is OPTIONAL and off by default (north_star lists
   sphere-vs-point-cloud only): when enabled, a configuration also collides if two spheres of a
   link pair in ``SELF_COLLISION_PAIRS`` (``TyGrit/rl/train.py:78-107``) touch,
   ``self_clearance = min_{(i,j)} (|c_i - c_j| - r_i - r_j)``, valid iff ``> 0`` (VAMP's own pair
   list is unavailable; the reference's RL reward uses exactly these link pairs).
6. Edge a->b over the 11-vector ``[x, y, theta, q0..q7]`` (no angle wrapping):
   ``n = max(ceil(|b-a|_2 * resolution / rake), 1)``, samples
   ``t_k = k / (rake n)``, ``k = 1 .. rake n`` (end included, start excluded),
   ``rake = 8``, ``resolution = 32``; edge valid iff every sample is valid.
   ``check_trajectory_collisions`` (``vamp_preview.py:306-322``) = any waypoint
   or any edge between consecutive waypoints collides.
"""

from __future__ import annotations

import numpy as np
from scipy.spatial import cKDTree

from . import fetch_model as fm
from . import fk

RAKE = 8
RESOLUTION = 32
POINT_RADIUS = 0.03
BAND = 1e-5


def quat_xyzw_to_matrix(q) -> np.ndarray:
    x, y, z, w = (float(v) for v in q)
    n = np.sqrt(x * x + y * y + z * z + w * w)
    x, y, z, w = x / n, y / n, z / n, w / n
    return np.array(
        [
            [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
            [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
            [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
        ]
    )


def robot_spheres_world(arm8, base_xyt, attachment=None):
    """Centres (B,S,3) and radii (S,) of robot (+ attached) spheres, fp64."""
    arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
    base = np.atleast_2d(np.asarray(base_xyt, dtype=np.float64))
    c = fk.sphere_centres(arm, base)
    r = fm.SPHERE_LOCAL[:, 3].copy()
    if attachment is not None and len(attachment["spheres"]) > 0:
        sph = np.asarray(attachment["spheres"], dtype=np.float64).reshape(-1, 4)
        Ro = quat_xyzw_to_matrix(attachment.get("quat", (0, 0, 0, 1)))
        to = np.asarray(attachment.get("pos", (0, 0, 0)), dtype=np.float64)
        local = sph[:, :3] @ Ro.T + to                       # in gripper_link frame
        q10 = np.concatenate([arm, np.zeros((arm.shape[0], 2))], axis=1)
        Rw, tw = fk.world_link_frames(q10, base)
        g = fm.LINK_INDEX["gripper_link"]
        ca = np.einsum("bij,kj->bki", Rw[:, g], local) + tw[:, g][:, None, :]
        c = np.concatenate([c, ca], axis=1)
        r = np.concatenate([r, sph[:, 3]])
    return c, r


class CloudOracle:
    """Exact nearest-neighbour clearance against a point cloud (+ sphere obstacles)."""

    def __init__(self, cloud_xyz, point_radius: float = POINT_RADIUS, obstacles=None):
        pts = np.asarray(cloud_xyz, dtype=np.float64).reshape(-1, 3)
        self.points = pts
        self.r_pt = float(point_radius)
        self.tree = cKDTree(pts) if len(pts) else None
        self.obstacles = (
            np.asarray(obstacles, dtype=np.float64).reshape(-1, 4)
            if obstacles is not None and len(obstacles)
            else np.zeros((0, 4))
        )

    def sphere_clearance(self, centres, radii, workers: int = -1) -> np.ndarray:
        """(..., S, 3), (S,) -> (..., S) signed clearance of each sphere (inf if no geometry)."""
        c = np.asarray(centres, dtype=np.float64)
        out = np.full(c.shape[:-1], np.inf)
        if self.tree is not None:
            d, _ = self.tree.query(c.reshape(-1, 3), k=1, workers=workers)
            out = d.reshape(c.shape[:-1]) - radii - self.r_pt
        for ox, oy, oz, orad in self.obstacles:
            d = np.linalg.norm(c - np.array([ox, oy, oz]), axis=-1) - radii - orad
            out = np.minimum(out, d)
        return out

    def config_clearance(self, arm8, base_xyt, attachment=None, chunk: int = 65536) -> np.ndarray:
        arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
        base = np.atleast_2d(np.asarray(base_xyt, dtype=np.float64))
        out = np.empty(arm.shape[0])
        for s in range(0, arm.shape[0], chunk):
            c, r = robot_spheres_world(arm[s : s + chunk], base[s : s + chunk], attachment)
            out[s : s + chunk] = self.sphere_clearance(c, r).min(axis=-1)
        return out

    def config_valid(self, arm8, base_xyt, attachment=None):
        """(valid (B,) bool, clearance (B,) f64)."""
        cl = self.config_clearance(arm8, base_xyt, attachment)
        return cl > 0.0, cl

    # ---- edges -----------------------------------------------------------
    def edge_clearance(self, arm0, base0, arm1, base1, attachment=None,
                       rake: int = RAKE, resolution: float = RESOLUTION):
        """Min clearance over the samples of each edge; returns (clear (E,), nsamp (E,))."""
        a, b = edge_vectors(arm0, base0), edge_vectors(arm1, base1)
        n = edge_n(a, b, rake, resolution)
        out = np.empty(len(a))
        for e in range(len(a)):
            q = edge_samples(a[e], b[e], int(n[e]), rake)
            out[e] = self.config_clearance(q[:, 3:], q[:, :3], attachment).min()
        return out, n * rake

    def edge_valid(self, arm0, base0, arm1, base1, attachment=None, **kw):
        cl, ns = self.edge_clearance(arm0, base0, arm1, base1, attachment, **kw)
        return cl > 0.0, cl, ns


def self_clearance(arm8) -> np.ndarray:
    """(B,) fp64 ``min_{(i,j) in pairs} (|c_i - c_j| - r_i - r_j)`` over the sphere pairs of
    ``fetch_model.SELF_COLLISION_PAIRS`` (``TyGrit/rl/train.py:78-107``); base pose irrelevant, head = 0."""
    arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
    out = np.empty(arm.shape[0])
    r = fm.SPHERE_LOCAL[:, 3]
    for s0 in range(0, arm.shape[0], 8192):
        c = fk.sphere_centres(arm[s0:s0 + 8192], np.zeros((min(8192, arm.shape[0] - s0), 3)))
        d = np.linalg.norm(c[:, fm.SELF_PAIR_I] - c[:, fm.SELF_PAIR_J], axis=-1) - r[fm.SELF_PAIR_I] - r[fm.SELF_PAIR_J]
        out[s0:s0 + 8192] = d.min(axis=1)
    return out


def edge_vectors(arm8, base_xyt) -> np.ndarray:
    """(E,11) fp64 rows ``[x, y, theta, q0..q7]``."""
    arm = np.atleast_2d(np.asarray(arm8, dtype=np.float64))
    base = np.atleast_2d(np.asarray(base_xyt, dtype=np.float64))
    return np.concatenate([base, arm], axis=1)


def edge_n(a11, b11, rake: int = RAKE, resolution: float = RESOLUTION) -> np.ndarray:
    """Number of rake blocks per edge: ``max(ceil(|b-a| * resolution / rake), 1)``.

    The squared norm is accumulated sequentially in fp64 in the order
    x, y, theta, q0..q7 (the CUDA kernel uses the same order in fp64 so the
    sample counts are identical).
    """
    d = np.asarray(b11, dtype=np.float64) - np.asarray(a11, dtype=np.float64)
    s = np.zeros(d.shape[0])
    for j in range(d.shape[1]):
        s = s + d[:, j] * d[:, j]
    return np.maximum(np.ceil(np.sqrt(s) * (resolution / rake)), 1).astype(np.int64)


def edge_samples(a11, b11, n: int, rake: int = RAKE) -> np.ndarray:
    """(rake*n, 11) sample configs ``a + (k / (rake n)) (b - a)``, k = 1..rake n."""
    k = np.arange(1, rake * n + 1, dtype=np.float64)[:, None]
    a = np.asarray(a11, dtype=np.float64)[None, :]
    b = np.asarray(b11, dtype=np.float64)[None, :]
    return a + (k / (rake * n)) * (b - a)
