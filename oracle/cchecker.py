"""ctypes wrapper of ``oracle/cpu_checker.c`` (ORACLE / CPU BASELINE — test infrastructure).

Used for full-size parity runs (fp64 exact clearance with OpenMP) and as the
timed CPU baseline (fp32 verdict path).  Never imported by ``tygrit_b200``.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import fetch_model as fm

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtg_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "cpu_checker.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared", "-o", _SO, src, "-lm"], check=True)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.tgo_cloud_build.restype = C.c_void_p
        L.tgo_cloud_build.argtypes = [C.c_void_p, C.c_int64, C.c_double, C.c_double]
        L.tgo_cloud_free.argtypes = [C.c_void_p]
        L.tgo_set_model.argtypes = [C.c_void_p, C.c_void_p]
        L.tgo_clearance_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_void_p]
        L.tgo_centres_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.tgo_check_configs_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int]
        L.tgo_check_edges_f32.restype = C.c_int64
        L.tgo_check_edges_f32.argtypes = [C.c_void_p] * 5 + [C.c_int64, C.c_int, C.c_double, C.c_void_p, C.c_int, C.c_int]
        L.tgo_set_self_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.tgo_self_clearance_f64.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.tgo_edge_clearance_f64.argtypes = [C.c_void_p] * 5 + [C.c_int64, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.tgo_threads.restype = C.c_int
        L.tgo_set_threads.argtypes = [C.c_int]
        xyzr = np.ascontiguousarray(fm.SPHERE_LOCAL, dtype=np.float64)
        link = np.ascontiguousarray(fm.SPHERE_LINK, dtype=np.int64)
        L.tgo_set_model(xyzr.ctypes.data, link.ctypes.data)
        pi = np.ascontiguousarray(fm.SELF_PAIR_I, dtype=np.int64)
        pj = np.ascontiguousarray(fm.SELF_PAIR_J, dtype=np.int64)
        L.tgo_set_self_pairs(pi.ctypes.data, pj.ctypes.data, pi.shape[0])
        _lib = L
    return _lib


def _f32(a, cols):
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] == cols, a.shape
    return a


class CCloud:
    """Uniform-grid cloud of the C oracle."""

    def __init__(self, cloud_xyz, point_radius: float = 0.03, cell: float = 0.1):
        self._L = lib()
        pts = np.ascontiguousarray(np.asarray(cloud_xyz, dtype=np.float32).reshape(-1, 3))
        self._h = C.c_void_p(self._L.tgo_cloud_build(pts.ctypes.data, pts.shape[0], float(point_radius), float(cell)))
        self.n = pts.shape[0]

    def __del__(self):  # pragma: no cover
        try:
            self._L.tgo_cloud_free(self._h)
        except Exception:
            pass

    def clearance(self, arm, base, cap: float = 0.25) -> np.ndarray:
        """fp64 exact clearance per config, clipped at ``cap`` (B,)."""
        arm, base = _f32(arm, 8), _f32(base, 3)
        out = np.empty(arm.shape[0], dtype=np.float64)
        self._L.tgo_clearance_f64(self._h, arm.ctypes.data, base.ctypes.data, arm.shape[0], float(cap), out.ctypes.data)
        return out

    def check_configs_f32(self, arm, base, threads: int = 0, self_collision: bool = False, naive: bool = False) -> np.ndarray:
        """fp32 verdict path (the timed CPU baseline): coarse-to-fine along the link chain; ``naive`` = all 111
        centres first, table order (the two forms give identical verdicts).  ``threads`` 0 = OpenMP default."""
        arm, base = _f32(arm, 8), _f32(base, 3)
        out = np.empty(arm.shape[0], dtype=np.uint8)
        self._L.tgo_check_configs_f32(self._h, arm.ctypes.data, base.ctypes.data, arm.shape[0], out.ctypes.data, int(threads),
                                      int(bool(self_collision)) | (2 if naive else 0))
        return out

    def check_edges_f32(self, arm0, base0, arm1, base1, rake: int = 8, resolution: float = 32.0, threads: int = 0,
                        self_collision: bool = False):
        a0, b0, a1, b1 = _f32(arm0, 8), _f32(base0, 3), _f32(arm1, 8), _f32(base1, 3)
        out = np.empty(a0.shape[0], dtype=np.uint8)
        n = self._L.tgo_check_edges_f32(self._h, a0.ctypes.data, b0.ctypes.data, a1.ctypes.data, b1.ctypes.data,
                                        a0.shape[0], int(rake), float(resolution), out.ctypes.data, int(threads),
                                        int(bool(self_collision)))
        return out, int(n)

    def edge_clearance(self, arm0, base0, arm1, base1, rake: int = 8, resolution: float = 32.0, cap: float = 0.25):
        a0, b0, a1, b1 = _f32(arm0, 8), _f32(base0, 3), _f32(arm1, 8), _f32(base1, 3)
        out = np.empty(a0.shape[0], dtype=np.float64)
        ns = np.empty(a0.shape[0], dtype=np.int64)
        self._L.tgo_edge_clearance_f64(self._h, a0.ctypes.data, b0.ctypes.data, a1.ctypes.data, b1.ctypes.data,
                                       a0.shape[0], int(rake), float(resolution), float(cap), out.ctypes.data, ns.ctypes.data)
        return out, ns


def centres_f64(arm, base, head=None) -> np.ndarray:
    arm, base = _f32(arm, 8), _f32(base, 3)
    h = _f32(head, 2) if head is not None else None
    out = np.empty((arm.shape[0], fm.N_SPHERES, 3), dtype=np.float64)
    lib().tgo_centres_f64(arm.ctypes.data, base.ctypes.data, h.ctypes.data if h is not None else None, arm.shape[0],
                          out.ctypes.data)
    return out


def self_clearance(arm) -> np.ndarray:
    """fp64 ``min (|c_i - c_j| - r_i - r_j)`` over the sphere pairs of SELF_COLLISION_PAIRS (rl/train.py:78-107), (B,)."""
    arm = _f32(arm, 8)
    out = np.empty(arm.shape[0], dtype=np.float64)
    lib().tgo_self_clearance_f64(arm.ctypes.data, arm.shape[0], out.ctypes.data)
    return out


def threads() -> int:
    return int(lib().tgo_threads())


def set_threads(n: int) -> int:
    """Set the OpenMP thread count of all later oracle calls; returns the count now in effect."""
    lib().tgo_set_threads(int(n))
    return threads()


def host_threads() -> int:
    """Host threads this process may use (``OMP_NUM_THREADS=1`` set by torchrun is deliberately ignored)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1
