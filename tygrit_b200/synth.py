"""Seeded synthetic workloads for tests and ``bench.py``.

Pure numpy so the oracle, the CPU baseline and the CUDA path all see the very
same bytes.  Seeds follow the reference's habit (``rng(42)`` for configurations,
``test/test_cross_validation.py:66-68``) and 1234 for clouds.

Nothing here is a product code path; it only manufactures inputs.
"""

from __future__ import annotations

import numpy as np

from .fetch_model import HEAD_LOWER, HEAD_UPPER, JOINT_LOWER, JOINT_UPPER

ROOM_HALF_X, ROOM_HALF_Y, ROOM_H = 4.0, 3.0, 2.6
GROUND_Z = 0.3  # belief_state/config.py:12 ground_z_threshold


# --------------------------------------------------------------------------- #
# clouds
# --------------------------------------------------------------------------- #
def _box_faces(lo, hi):
    """Six rectangles (origin, edge_u, edge_v) of an axis-aligned box."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    d = hi - lo
    faces = []
    for ax in range(3):
        u, v = (ax + 1) % 3, (ax + 2) % 3
        eu, ev = np.zeros(3), np.zeros(3)
        eu[u], ev[v] = d[u], d[v]
        for side in (lo, hi):
            o = lo.copy()
            o[ax] = side[ax]
            faces.append((o, eu, ev))
    return faces


def _sample_rects(rng, rects, n):
    area = np.array([np.linalg.norm(np.cross(eu, ev)) for _, eu, ev in rects])
    which = rng.choice(len(rects), size=n, p=area / area.sum())
    a, b = rng.random(n), rng.random(n)
    o = np.array([r[0] for r in rects])[which]
    eu = np.array([r[1] for r in rects])[which]
    ev = np.array([r[2] for r in rects])[which]
    return o + a[:, None] * eu + b[:, None] * ev


def room_boxes(seed: int = 1234):
    """Furniture cuboids + two table slabs of the 'ReplicaCAD-scale' room (lo, hi) pairs."""
    rng = np.random.default_rng(seed)
    boxes = []
    n_furn = int(rng.integers(10, 21))
    for _ in range(n_furn):
        sx, sy = rng.uniform(0.4, 1.5, 2)
        h = rng.uniform(0.4, 1.8)
        cx = rng.uniform(-ROOM_HALF_X + sx / 2, ROOM_HALF_X - sx / 2)
        cy = rng.uniform(-ROOM_HALF_Y + sy / 2, ROOM_HALF_Y - sy / 2)
        boxes.append(((cx - sx / 2, cy - sy / 2, 0.0), (cx + sx / 2, cy + sy / 2, h)))
    for _ in range(2):  # table tops: thin slabs at 0.72 m
        sx, sy = rng.uniform(0.8, 1.8, 2)
        cx = rng.uniform(-ROOM_HALF_X + sx / 2, ROOM_HALF_X - sx / 2)
        cy = rng.uniform(-ROOM_HALF_Y + sy / 2, ROOM_HALF_Y - sy / 2)
        boxes.append(((cx - sx / 2, cy - sy / 2, 0.69), (cx + sx / 2, cy + sy / 2, 0.72)))
    return boxes


def room_cloud(n_points: int, seed: int = 1234, perturb_step: int | None = None) -> np.ndarray:
    """(n_points, 3) f32 room cloud: 4 walls + furniture + table tops, ground removed.

    ``perturb_step`` (C5): furniture boxes 0 and 1 are translated by a
    step-dependent offset so ~10 % of the cloud moves between scheduler steps.
    """
    boxes = room_boxes(seed)
    if perturb_step is not None:
        prng = np.random.default_rng(seed + 7919 * (perturb_step + 1))
        for i in (0, 1):
            lo, hi = np.array(boxes[i][0]), np.array(boxes[i][1])
            sh = np.append(prng.uniform(-0.3, 0.3, 2), 0.0)
            boxes[i] = (tuple(lo + sh), tuple(hi + sh))
    rects = []
    X, Y, H = ROOM_HALF_X, ROOM_HALF_Y, ROOM_H
    rects += [
        (np.array([-X, -Y, 0.0]), np.array([2 * X, 0, 0.0]), np.array([0, 0, H])),
        (np.array([-X, Y, 0.0]), np.array([2 * X, 0, 0.0]), np.array([0, 0, H])),
        (np.array([-X, -Y, 0.0]), np.array([0, 2 * Y, 0.0]), np.array([0, 0, H])),
        (np.array([X, -Y, 0.0]), np.array([0, 2 * Y, 0.0]), np.array([0, 0, H])),
    ]
    for lo, hi in boxes:
        rects += _box_faces(lo, hi)
    rng = np.random.default_rng(seed + 1)
    out = np.empty((0, 3))
    while out.shape[0] < n_points:  # oversample, drop ground, repeat until full
        p = _sample_rects(rng, rects, int((n_points - out.shape[0]) * 1.3) + 64)
        p = p + rng.normal(0.0, 0.002, p.shape)
        out = np.vstack([out, p[p[:, 2] > GROUND_Z]])
    return np.ascontiguousarray(out[:n_points], dtype=np.float32)


def tabletop_cloud(n_points: int = 50_000, seed: int = 1234) -> np.ndarray:
    """(n_points,3) f32: table top 1.0 x 1.6 m at (0.9, 0, 0.72) + 12 clutter boxes."""
    rng = np.random.default_rng(seed)
    n_top = n_points // 2
    top = np.stack(
        [rng.uniform(0.4, 1.4, n_top), rng.uniform(-0.8, 0.8, n_top), 0.72 + rng.normal(0, 0.002, n_top)],
        axis=1,
    )
    rects = []
    for _ in range(12):
        e = rng.uniform(0.05, 0.25, 3)
        cx, cy = rng.uniform(0.5, 1.3), rng.uniform(-0.7, 0.7)
        rects += _box_faces((cx - e[0] / 2, cy - e[1] / 2, 0.72), (cx + e[0] / 2, cy + e[1] / 2, 0.72 + e[2]))
    clutter = _sample_rects(rng, rects, n_points - n_top) + rng.normal(0, 0.002, (n_points - n_top, 3))
    return np.ascontiguousarray(np.vstack([top, clutter]), dtype=np.float32)


# --------------------------------------------------------------------------- #
# configurations / edges / robot states
# --------------------------------------------------------------------------- #
def random_configs(n: int, seed: int = 42, base_box=(-3.5, 3.5, -2.5, 2.5)):
    """arm (n,8) f32 ~ U(limits), base (n,3) f32 [x, y, theta] over the floor rectangle."""
    rng = np.random.default_rng(seed)
    arm = rng.uniform(JOINT_LOWER, JOINT_UPPER, (n, 8))
    x0, x1, y0, y1 = base_box
    base = np.stack([rng.uniform(x0, x1, n), rng.uniform(y0, y1, n), rng.uniform(-np.pi, np.pi, n)], axis=1)
    return arm.astype(np.float32), base.astype(np.float32)


def random_edges(n: int, seed: int = 42, base_box=(-3.5, 3.5, -2.5, 2.5), rrt_like: bool = False):
    """Arm-only edges at a random base (as ``plan_arm_rrtc``): (arm0, base0, arm1, base1) f32.

    ``rrt_like``: ``b = a + 0.5 * unit`` (|delta| = 0.5 -> 16 samples/edge) instead of
    i.i.d. endpoints (mean ~171 samples/edge).
    """
    rng = np.random.default_rng(seed)
    a = rng.uniform(JOINT_LOWER, JOINT_UPPER, (n, 8))
    if rrt_like:
        u = rng.normal(size=(n, 8))
        b = a + 0.5 * u / np.linalg.norm(u, axis=1, keepdims=True)
    else:
        b = rng.uniform(JOINT_LOWER, JOINT_UPPER, (n, 8))
    x0, x1, y0, y1 = base_box
    base = np.stack([rng.uniform(x0, x1, n), rng.uniform(y0, y1, n), rng.uniform(-np.pi, np.pi, n)], axis=1)
    f = np.float32
    return a.astype(f), base.astype(f), b.astype(f), base.astype(f).copy()


def random_states(n: int, seed: int = 42, base_box=(-3.0, 3.0, -2.0, 2.0)) -> np.ndarray:
    """(n,13) f32 robot states ``[x, y, theta, q0..q7, head_pan, head_tilt]``."""
    rng = np.random.default_rng(seed)
    arm = rng.uniform(JOINT_LOWER, JOINT_UPPER, (n, 8))
    head = rng.uniform(HEAD_LOWER, HEAD_UPPER, (n, 2))
    x0, x1, y0, y1 = base_box
    base = np.stack([rng.uniform(x0, x1, n), rng.uniform(y0, y1, n), rng.uniform(-np.pi, np.pi, n)], axis=1)
    return np.concatenate([base, arm, head], axis=1).astype(np.float32)


HEAD_CAMERA_K = np.array([[154.10223, 0.0, 320.0], [0.0, 154.10223, 240.0], [0.0, 0.0, 1.0]])


# --------------------------------------------------------------------------- #
# device-side generators (bench.py: 64 Mi configurations never cross PCIe)
# --------------------------------------------------------------------------- #
TUCK_ARM = (0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0)   # the reference tests' tuck configuration (test/test_planner.py:35)


def device_configs(n: int, seed: int, device, base_box=(-3.5, 3.5, -2.5, 2.5)):
    """arm (n,8), base (n,3) f32 CUDA tensors: the distribution of :func:`random_configs`, torch Philox stream."""
    import torch

    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    lo = torch.tensor(JOINT_LOWER, dtype=torch.float32, device=device)
    hi = torch.tensor(JOINT_UPPER, dtype=torch.float32, device=device)
    arm = torch.rand(n, 8, generator=gen, device=device) * (hi - lo) + lo
    blo = torch.tensor([base_box[0], base_box[2], -np.pi], dtype=torch.float32, device=device)
    bhi = torch.tensor([base_box[1], base_box[3], np.pi], dtype=torch.float32, device=device)
    base = torch.rand(n, 3, generator=gen, device=device) * (bhi - blo) + blo
    return arm, base


def free_base_configs(world, n: int, seed: int, device, base_box=(-3.5, 3.5, -2.5, 2.5), chunk: int = 1 << 24):
    """Planner-like ("free-heavy") batch: the base is drawn uniformly from the floor positions at which the robot with
    its arm TUCKED is collision-free in ``world`` (rejection sampling with the engine itself), the arm uniformly from
    the joint limits.  Returns (arm, base, accepted fraction of the floor)."""
    import torch

    gen = torch.Generator(device=device)
    gen.manual_seed(seed + 7919)
    blo = torch.tensor([base_box[0], base_box[2], -np.pi], dtype=torch.float32, device=device)
    bhi = torch.tensor([base_box[1], base_box[3], np.pi], dtype=torch.float32, device=device)
    tuck = torch.tensor(TUCK_ARM, dtype=torch.float32, device=device)
    out, have, tried, kept = [], 0, 0, 0
    while have < n:
        m = int(min(chunk, max(4096, 3 * (n - have))))
        cand = torch.rand(m, 3, generator=gen, device=device) * (bhi - blo) + blo
        ok = world.check_configs(tuck.expand(m, 8).contiguous(), cand).bool()
        good = cand[ok]
        tried, kept = tried + m, kept + int(good.shape[0])
        if kept == 0 and tried >= 4 * chunk:
            raise RuntimeError("free_base_configs: no free floor position found")
        good = good[: n - have]
        out.append(good)
        have += int(good.shape[0])
    base = torch.cat(out).contiguous()
    lo = torch.tensor(JOINT_LOWER, dtype=torch.float32, device=device)
    hi = torch.tensor(JOINT_UPPER, dtype=torch.float32, device=device)
    arm = torch.rand(n, 8, generator=gen, device=device) * (hi - lo) + lo
    return arm, base, kept / max(tried, 1)
