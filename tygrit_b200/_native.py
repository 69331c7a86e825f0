"""ctypes binding of ``libtygrit_b200.so`` (the C ABI in ``include/tygrit_b200.h``).

There is no CPU fallback: if the library is missing or a call fails, a
``RuntimeError`` is raised.  PyTorch is used only for device memory and streams.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TYGRIT_B200_LIB") or os.path.join(_HERE, "lib", "libtygrit_b200.so")

TG_N_SPHERES = 111
TG_N_LINKS = 15
TG_MAX_OBSTACLES = 64
TG_MAX_ATTACHED = 64
TG_NOALLOC_MAX_CONFIGS = 262144  # configurations per tg_check_configs_noalloc call (static scratch)
TG_E_CAPACITY = -3

_lib = None

c_f32p = C.c_void_p  # device or host float*; passed as raw addresses
_SIGS = {
    "tg_last_error": (C.c_char_p, []),
    "tg_version": (C.c_int, []),
    "tg_env_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_float, C.c_float, C.c_int64]),
    "tg_env_destroy": (C.c_int, [C.c_void_p]),
    "tg_env_build": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "tg_env_build_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "tg_env_set_bounds": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "tg_env_set_point_radius": (C.c_int, [C.c_void_p, C.c_float]),
    "tg_env_set_cells_per_metre": (C.c_int, [C.c_void_p, C.c_float]),
    "tg_env_set_cell_shape": (C.c_int, [C.c_void_p, C.c_float, C.c_float]),
    "tg_env_set_obstacles": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "tg_env_set_attachment": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "tg_env_set_self_collision": (C.c_int, [C.c_void_p, C.c_int]),
    "tg_env_export": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "tg_env_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "tg_env_blob_capacity": (C.c_size_t, [C.c_void_p]),
    "tg_env_blob_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "tg_env_grid_info": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tg_env_stats": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tg_env_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "tg_env_set_coarse": (C.c_int, [C.c_void_p, C.c_int]),
    "tg_env_set_queue_limit": (C.c_int, [C.c_void_p, C.c_int64]),
    "tg_env_timing_read": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tg_env_counters": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "tg_debug_stats": (C.c_int, [C.c_void_p, C.c_int]),
    "tg_fk_spheres": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_fk_frames": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_check_configs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "tg_check_configs_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_check_edges": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p,
         C.c_void_p, C.c_void_p],
    ),
    "tg_pack_bits": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_unpack_bits": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_check_configs_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_check_edges_host": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p,
         C.c_void_p],
    ),
    "tg_filter_points": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "tg_self_filter_depth": (
        C.c_int,
        [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_float,
         C.c_void_p, C.c_void_p, C.c_void_p],
    ),
    "tg_compact_points": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_voxel_first": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "tg_ingest_depth": (
        C.c_int,
        [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float,
         C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p],
    ),
    "tg_carve_flags": (
        C.c_int,
        [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p,
         C.c_float, C.c_void_p, C.c_void_p],
    ),
    "tg_fk_frames_f64": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "tg_check_occlusion_frames": (
        C.c_int,
        [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    ),
    "tg_check_occlusion": (
        C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    ),
    "tg_first_valid": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]),
    "tg_planner_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_void_p, C.c_int, C.c_int]),
    "tg_planner_destroy": (C.c_int, [C.c_void_p]),
    "tg_planner_rrt_connect": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_uint32, C.c_int, C.c_float, C.c_double,
         C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p],
    ),
    "tg_planner_shortcut": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "tg_planner_base_path": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_int,
         C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p],
    ),
    "tg_planner_fcit": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_uint32, C.c_int,
         C.c_float, C.c_double, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p],
    ),
    "tg_c5_run": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_int, C.c_float,
         C.c_void_p],
    ),
    "tg_synth_depth": (
        C.c_int,
        [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_uint64, C.c_void_p,
         C.c_void_p],
    ),
}

EXPORTED_SYMBOLS = tuple(_SIGS.keys())


def lib() -> C.CDLL:
    """Load the CUDA library (once).  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"tygrit_b200: {LIB_PATH} is missing — run `python -m tygrit_b200.build` "
                "(there is no CPU fallback for the collision path)"
            )
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(handle, name)  # AttributeError = ABI mismatch: let it surface
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().tg_last_error()
        raise RuntimeError(f"tygrit_b200: {what} failed ({rc}): {msg.decode() if msg else ''}")
