"""``CollisionWorld`` — thin Python owner of one ``tg_env`` (device buffers of a collision world).

Tensor-in / tensor-out batch surface over the C ABI; every method enqueues on
``torch.cuda.current_stream()`` and returns without synchronising unless it
says otherwise.  Host (numpy) variants go through the library's pinned,
double-buffered pipeline.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as N


def _stream_ptr(device: torch.device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


def _f32c(t: torch.Tensor, device: torch.device, cols: int, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(np.asarray(t, dtype=np.float32))
    t = t.to(device=device, dtype=torch.float32).contiguous()
    if t.ndim != 2 or t.shape[1] != cols:
        raise ValueError(f"{name}: expected shape (N, {cols}), got {tuple(t.shape)}")
    return t


class CollisionWorld:
    """One collision world on one GPU: cloud structure + obstacles + attachment."""

    def __init__(
        self,
        device: int | str | torch.device = 0,
        max_points: int = 1 << 20,
        voxels_per_metre: float = 50.0,
        point_radius: float = 0.03,
        max_voxels: int = 0,
    ) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("tygrit_b200: no CUDA device — the collision engine has no CPU fallback")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        torch.cuda.init()
        with torch.cuda.device(self.device):
            torch.zeros(1, device=self.device)  # make sure the primary context exists
        self._lib = N.lib()
        h = C.c_void_p()
        N.check(
            self._lib.tg_env_create(C.byref(h), self.device.index, int(max_points), float(voxels_per_metre),
                                    float(point_radius), int(max_voxels)),
            "tg_env_create",
        )
        self._h = h
        self.max_points = int(max_points)
        self.n_points = 0

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.tg_env_destroy(self._h)
            self._h = None

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ---- world ------------------------------------------------------------
    def set_point_radius(self, r: float) -> None:
        N.check(self._lib.tg_env_set_point_radius(self._h, float(r)), "tg_env_set_point_radius")

    def set_cells_per_metre(self, cpm: float) -> None:
        """Resolution of the narrow-phase cell grid, used by the next :meth:`build`."""
        N.check(self._lib.tg_env_set_cells_per_metre(self._h, float(cpm)), "tg_env_set_cells_per_metre")

    def set_cell_shape(self, cpm_x: float, cpm_yz: float) -> None:
        """Anisotropic narrow-phase cells (x = the row direction), used by the next :meth:`build`."""
        N.check(self._lib.tg_env_set_cell_shape(self._h, float(cpm_x), float(cpm_yz)), "tg_env_set_cell_shape")

    def set_bounds(self, lo=None, hi=None) -> None:
        if lo is None or hi is None:
            N.check(self._lib.tg_env_set_bounds(self._h, None, None), "tg_env_set_bounds")
            return
        lo_a = (C.c_float * 3)(*[float(v) for v in lo])
        hi_a = (C.c_float * 3)(*[float(v) for v in hi])
        N.check(self._lib.tg_env_set_bounds(self._h, lo_a, hi_a), "tg_env_set_bounds")

    def build(self, points) -> None:
        """Replace the cloud.  ``points``: (N,3) CUDA tensor, or numpy/CPU tensor (host path)."""
        if isinstance(points, torch.Tensor) and points.is_cuda:
            p = _f32c(points, self.device, 3, "points")
            with torch.cuda.device(self.device):
                N.check(self._lib.tg_env_build(self._h, p.data_ptr(), p.shape[0], _stream_ptr(self.device)), "tg_env_build")
            self._keep = p  # the build kernels read it asynchronously
        else:
            a = np.ascontiguousarray(np.asarray(points, dtype=np.float32).reshape(-1, 3))
            with torch.cuda.device(self.device):
                N.check(
                    self._lib.tg_env_build_host(self._h, a.ctypes.data, a.shape[0], _stream_ptr(self.device)),
                    "tg_env_build_host",
                )
                torch.cuda.current_stream(self.device).synchronize()  # `a` may be freed by the caller
        self.n_points = int(points.shape[0]) if hasattr(points, "shape") else len(points)

    def set_obstacles(self, xyzr) -> None:
        a = np.ascontiguousarray(np.asarray(xyzr, dtype=np.float32).reshape(-1, 4))
        N.check(self._lib.tg_env_set_obstacles(self._h, a.ctypes.data, a.shape[0], _stream_ptr(self.device)),
                "tg_env_set_obstacles")

    def set_attachment(self, xyzr, offset_pos=None, offset_quat=None) -> None:
        a = np.ascontiguousarray(np.asarray(xyzr, dtype=np.float32).reshape(-1, 4))
        pos = (C.c_float * 3)(*[float(v) for v in (offset_pos if offset_pos is not None else (0, 0, 0))])
        quat = (C.c_float * 4)(*[float(v) for v in (offset_quat if offset_quat is not None else (0, 0, 0, 1))])
        N.check(
            self._lib.tg_env_set_attachment(self._h, a.ctypes.data, a.shape[0], pos, quat, _stream_ptr(self.device)),
            "tg_env_set_attachment",
        )

    def set_self_collision(self, on: bool) -> None:
        """Optional sphere-vs-sphere term over the reference's SELF_COLLISION_PAIRS (rl/train.py:78-107); off by default."""
        N.check(self._lib.tg_env_set_self_collision(self._h, int(bool(on))), "tg_env_set_self_collision")

    def stats(self) -> dict[str, float]:
        out = (C.c_double * 8)()
        N.check(self._lib.tg_env_stats(self._h, out), "tg_env_stats")
        k = ("n_points", "n_dropped", "nx", "ny", "nz", "cells", "voxels_per_metre", "slack")
        return dict(zip(k, [float(v) for v in out]))

    def grid_info(self) -> dict[str, float]:
        """Layout of the built structure inside the blob (tests / tools)."""
        out = (C.c_double * 20)()
        N.check(self._lib.tg_env_grid_info(self._h, out), "tg_env_grid_info")
        k = ("empty", "nx", "ny", "nz", "bias_x", "bias_y", "bias_z", "voxels_per_metre", "q", "slack", "window", "off_df", "off_cg",
             "gnx", "gny", "gnz", "cg_r", "cg_q")
        return dict(zip(k, [float(v) for v in out]))

    def counters(self) -> dict[str, int]:
        """Counters since the last call (synchronises the current stream)."""
        out = (C.c_int64 * 4)()
        N.check(self._lib.tg_env_counters(self._h, out, _stream_ptr(self.device)), "tg_env_counters")
        k = ("definite_collisions", "narrow_items", "overflow_inline", "launches")
        return dict(zip(k, [int(v) for v in out]))

    def set_queue_limit(self, items: int) -> None:
        """Test hook: shrink the ambiguous queue (0 = whole queue) to exercise the inline overflow path."""
        N.check(self._lib.tg_env_set_queue_limit(self._h, int(items)), "tg_env_set_queue_limit")

    def set_coarse(self, on: bool) -> None:
        """Test / tuning hook: switch the shared-memory coarse far-field variant of the configuration kernel on or off."""
        N.check(self._lib.tg_env_set_coarse(self._h, int(bool(on))), "tg_env_set_coarse")

    def set_timing(self, on: bool) -> None:
        N.check(self._lib.tg_env_set_timing(self._h, int(bool(on))), "tg_env_set_timing")

    def timing_read(self) -> dict[str, float]:
        """Device time of the check-config launches since the last read (waits for them)."""
        out = (C.c_double * 3)()
        N.check(self._lib.tg_env_timing_read(self._h, out), "tg_env_timing_read")
        return {"phase1_ms": float(out[0]), "phase2_ms": float(out[1]), "launch_pairs": int(out[2])}

    def blob_view(self, nbytes: int | None = None) -> torch.Tensor:
        """Zero-copy uint8 CUDA view of the env's structure blob (first ``nbytes`` bytes, default: capacity).

        Rank 0 broadcasts ``blob_view(used_bytes)``; the other ranks receive straight into their own
        ``blob_view(used_bytes)`` and then call :meth:`adopt_blob`.
        """
        cap = int(self._lib.tg_env_blob_capacity(self._h))
        n = cap if nbytes is None else int(nbytes)
        if n > cap:
            raise ValueError(f"blob_view: {n} bytes exceed the env capacity {cap}")
        ptr = C.c_void_p()
        N.check(self._lib.tg_env_blob_ptr(self._h, C.byref(ptr)), "tg_env_blob_ptr")

        class _Holder:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (int(ptr.value), False), "version": 2}

        return torch.as_tensor(_Holder(), device=self.device)

    def used_bytes(self) -> int:
        ptr, nbytes = C.c_void_p(), C.c_size_t()
        N.check(self._lib.tg_env_export(self._h, C.byref(ptr), C.byref(nbytes)), "tg_env_export")
        return int(nbytes.value)

    def adopt_blob(self, nbytes: int) -> None:
        """Adopt a structure that was broadcast into ``blob_view(nbytes)`` (synchronises once)."""
        v = self.blob_view(nbytes)
        with torch.cuda.device(self.device):
            N.check(self._lib.tg_env_import(self._h, v.data_ptr(), int(nbytes), _stream_ptr(self.device)), "tg_env_import")

    def import_blob(self, blob: torch.Tensor) -> None:
        """Adopt a structure held in another uint8 CUDA tensor (copied into the env)."""
        if not blob.is_cuda or blob.dtype != torch.uint8:
            raise ValueError("import_blob: expected a uint8 CUDA tensor")
        blob = blob.contiguous()
        with torch.cuda.device(self.device):
            N.check(self._lib.tg_env_import(self._h, blob.data_ptr(), blob.numel(), _stream_ptr(self.device)), "tg_env_import")

    # ---- verdicts ----------------------------------------------------------
    def check_configs(self, arm, base, clearance: bool = False):
        """(B,8), (B,3) CUDA f32 -> valid (B,) uint8 CUDA [, clearance (B,) f32]."""
        arm = _f32c(arm, self.device, 8, "arm")
        base = _f32c(base, self.device, 3, "base")
        if arm.shape[0] != base.shape[0]:
            raise ValueError("arm and base must have the same number of rows")
        B = arm.shape[0]
        valid = torch.empty(B, dtype=torch.uint8, device=self.device)
        cl = torch.empty(B, dtype=torch.float32, device=self.device) if clearance else None
        with torch.cuda.device(self.device):
            N.check(
                self._lib.tg_check_configs(self._h, arm.data_ptr(), base.data_ptr(), B, valid.data_ptr(),
                                           cl.data_ptr() if cl is not None else None, _stream_ptr(self.device)),
                "tg_check_configs",
            )
        return (valid, cl) if clearance else valid

    def check_configs_into(self, arm: torch.Tensor, base: torch.Tensor, valid: torch.Tensor) -> None:
        """Zero-allocation variant for benchmarking loops (inputs must already be CUDA f32 contiguous)."""
        with torch.cuda.device(self.device):
            N.check(
                self._lib.tg_check_configs(self._h, arm.data_ptr(), base.data_ptr(), arm.shape[0], valid.data_ptr(), None,
                                           _stream_ptr(self.device)),
                "tg_check_configs",
            )

    def check_edges(self, arm0, base0, arm1, base1, rake: int = 8, resolution: float = 32.0, count: bool = False):
        a0, b0 = _f32c(arm0, self.device, 8, "arm0"), _f32c(base0, self.device, 3, "base0")
        a1, b1 = _f32c(arm1, self.device, 8, "arm1"), _f32c(base1, self.device, 3, "base1")
        E = a0.shape[0]
        if not (b0.shape[0] == a1.shape[0] == b1.shape[0] == E):
            raise ValueError("edge arrays must have the same number of rows")
        valid = torch.empty(E, dtype=torch.uint8, device=self.device)
        nchk = torch.zeros(1, dtype=torch.int64, device=self.device) if count else None
        with torch.cuda.device(self.device):
            N.check(
                self._lib.tg_check_edges(self._h, a0.data_ptr(), b0.data_ptr(), a1.data_ptr(), b1.data_ptr(), E, int(rake),
                                         float(resolution), valid.data_ptr(),
                                         nchk.data_ptr() if nchk is not None else None, _stream_ptr(self.device)),
                "tg_check_edges",
            )
        return (valid, nchk) if count else valid

    def run_replanning_steps(self, clouds, n_edges: int, steps: int = 100, warmup: int = 3, seed: int = 42,
                             base_box=(-3.5, 3.5, -2.5, 2.5), rake: int = 8, resolution: float = 32.0) -> dict[str, float]:
        """BASELINE config 5 as CUDA-graph replays driven from C (``tg_c5_run``): per step fresh device-generated edges,
        a full rebuild on ``clouds[i % len(clouds)]`` (CUDA (N,3) f32 tensors) and ``n_edges`` edge validations.
        Needs fixed bounds (:meth:`set_bounds`).  Synchronous."""
        cl = [_f32c(c, self.device, 3, "cloud") for c in clouds]
        ptrs = (C.c_void_p * len(cl))(*[c.data_ptr() for c in cl])
        ns = (C.c_int64 * len(cl))(*[c.shape[0] for c in cl])
        box = (C.c_float * 4)(*[float(v) for v in base_box])
        out = (C.c_double * 8)()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize(self.device)
            N.check(self._lib.tg_c5_run(self._h, ptrs, ns, len(cl), int(n_edges), int(steps), int(warmup), int(seed), box, int(rake),
                                        float(resolution), out), "tg_c5_run")
        k = ("ms_per_step", "build_ms", "valid_fraction", "config_checks_per_step", "graphs", "nodes_per_step_graph", "min_step_ms",
             "max_step_ms")
        return dict(zip(k, [float(v) for v in out]))

    # ---- host (numpy) variants: the reference-facing path -------------------
    def check_configs_host(self, arm: np.ndarray, base: np.ndarray, out: np.ndarray | None = None,
                           clearance: bool = False):
        arm = np.ascontiguousarray(arm, dtype=np.float32)
        base = np.ascontiguousarray(base, dtype=np.float32)
        if arm.ndim != 2 or arm.shape[1] != 8 or base.shape != (arm.shape[0], 3):
            raise ValueError("expected arm (B,8) and base (B,3)")
        B = arm.shape[0]
        valid = out if out is not None else np.empty(B, dtype=np.uint8)
        cl = np.empty(B, dtype=np.float32) if clearance else None
        N.check(
            self._lib.tg_check_configs_host(self._h, arm.ctypes.data, base.ctypes.data, B, valid.ctypes.data,
                                            cl.ctypes.data if cl is not None else None),
            "tg_check_configs_host",
        )
        return (valid, cl) if clearance else valid

    def check_edges_host(self, arm0, base0, arm1, base1, rake: int = 8, resolution: float = 32.0,
                         out: np.ndarray | None = None):
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (arm0, base0, arm1, base1)]
        E = arrs[0].shape[0]
        valid = out if out is not None else np.empty(E, dtype=np.uint8)
        n = C.c_int64(0)
        N.check(
            self._lib.tg_check_edges_host(self._h, arrs[0].ctypes.data, arrs[1].ctypes.data, arrs[2].ctypes.data,
                                          arrs[3].ctypes.data, E, int(rake), float(resolution), valid.ctypes.data,
                                          C.byref(n)),
            "tg_check_edges_host",
        )
        return valid, int(n.value)


# ---- stateless kernels -------------------------------------------------------
def fk_spheres(arm, base, head=None) -> torch.Tensor:
    """World centres of the 111 spheres, (B,111,3) CUDA f32, reference sphere order."""
    dev = arm.device if isinstance(arm, torch.Tensor) and arm.is_cuda else torch.device("cuda", torch.cuda.current_device())
    arm = _f32c(arm, dev, 8, "arm")
    base = _f32c(base, dev, 3, "base")
    h = _f32c(head, dev, 2, "head") if head is not None else None
    out = torch.empty(arm.shape[0], N.TG_N_SPHERES, 3, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        N.check(
            N.lib().tg_fk_spheres(arm.data_ptr(), base.data_ptr(), h.data_ptr() if h is not None else None, arm.shape[0],
                                  out.data_ptr(), _stream_ptr(dev)),
            "tg_fk_spheres",
        )
    return out


def pack_verdicts(valid: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """(n,) uint8 CUDA verdict bytes -> (ceil(n/8),) uint8 bit mask (bit i & 7 of byte i // 8)."""
    n = valid.numel()
    if out is None:
        out = torch.empty((n + 7) // 8, dtype=torch.uint8, device=valid.device)
    with torch.cuda.device(valid.device):
        N.check(N.lib().tg_pack_bits(valid.data_ptr(), n, out.data_ptr(), _stream_ptr(valid.device)), "tg_pack_bits")
    return out


def unpack_verdicts(bits: torch.Tensor, n: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """Inverse of :func:`pack_verdicts`."""
    if out is None:
        out = torch.empty(n, dtype=torch.uint8, device=bits.device)
    with torch.cuda.device(bits.device):
        N.check(N.lib().tg_unpack_bits(bits.data_ptr(), int(n), out.data_ptr(), _stream_ptr(bits.device)), "tg_unpack_bits")
    return out


def fk_frames(q10) -> torch.Tensor:
    """Link frames in base_link, (B,15,4,4) CUDA f32."""
    dev = q10.device if isinstance(q10, torch.Tensor) and q10.is_cuda else torch.device("cuda", torch.cuda.current_device())
    q = _f32c(q10, dev, 10, "q10")
    out = torch.empty(q.shape[0], N.TG_N_LINKS, 4, 4, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        N.check(N.lib().tg_fk_frames(q.data_ptr(), q.shape[0], out.data_ptr(), _stream_ptr(dev)), "tg_fk_frames")
    return out


def filter_points_mask(points: torch.Tensor, centres: torch.Tensor, radius: float) -> torch.Tensor:
    """keep mask (N,) uint8 CUDA: 1 iff the point is outside every sphere of ``radius``."""
    dev = points.device
    p = _f32c(points, dev, 3, "points")
    c = _f32c(centres, dev, 3, "centres")
    keep = torch.empty(p.shape[0], dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        N.check(
            N.lib().tg_filter_points(p.data_ptr(), p.shape[0], c.data_ptr(), c.shape[0], float(radius), keep.data_ptr(),
                                     _stream_ptr(dev)),
            "tg_filter_points",
        )
    return keep


def self_filter_depth(depth: torch.Tensor, K, states: torch.Tensor, radius: float = 0.08, z_range=(0.2, 3.0),
                      stride: int = 1, want_world: bool = False, mask: torch.Tensor | None = None):
    """Batched depth self-filter: depth (S,H,W), states (S,13) -> mask (S,H',W') uint8 [, world (S,H',W',3)]."""
    dev = depth.device
    if depth.ndim != 3 or depth.dtype != torch.float32 or not depth.is_contiguous():
        raise ValueError("depth must be a contiguous (S,H,W) float32 CUDA tensor")
    st = _f32c(states, dev, 13, "states")
    S, H, W = depth.shape
    if st.shape[0] != S:
        raise ValueError("states must have one row per depth frame")
    Hs, Ws = (H + stride - 1) // stride, (W + stride - 1) // stride
    if mask is None:
        mask = torch.empty(S, Hs, Ws, dtype=torch.uint8, device=dev)
    world = torch.empty(S, Hs, Ws, 3, dtype=torch.float32, device=dev) if want_world else None
    Kd = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(3, 3))
    with torch.cuda.device(dev):
        N.check(
            N.lib().tg_self_filter_depth(depth.data_ptr(), H, W, int(stride), Kd.ctypes.data, st.data_ptr(), S,
                                         float(radius), float(z_range[0]), float(z_range[1]), mask.data_ptr(),
                                         world.data_ptr() if world is not None else None, _stream_ptr(dev)),
            "tg_self_filter_depth",
        )
    return (mask, world) if want_world else mask
