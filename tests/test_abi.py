"""CPU: the C-ABI library builds, loads and exports every symbol include/tygrit_b200.h declares;
host-side logic (types, configs, factories, Halton, interpolation) behaves like the reference's."""

from __future__ import annotations

import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols() -> list[str]:
    txt = open(os.path.join(ROOT, "include", "tygrit_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(tg_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built_lib):
    lib_path, oracle_path = built_lib
    assert os.path.exists(lib_path)
    h = ctypes.CDLL(lib_path)          # loads without a GPU (no compute call is made)
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(h, n), f"{n} declared in include/tygrit_b200.h but not exported"
    h.tg_version.restype = ctypes.c_int
    assert h.tg_version() >= 100


def test_python_binding_covers_header(built_lib):
    from tygrit_b200 import _native

    assert sorted(_native.EXPORTED_SYMBOLS) == _declared_symbols()
    _native.lib()                      # sets argtypes on every symbol; AttributeError = ABI drift


def test_argument_errors_without_gpu(built_lib):
    from tygrit_b200 import _native

    L = _native.lib()
    assert L.tg_env_create(None, 0, 1000, 50.0, 0.03, 0) == -1          # TG_E_ARG: out is NULL
    assert b"out is NULL" in L.tg_last_error()
    h = ctypes.c_void_p()
    assert L.tg_env_create(ctypes.byref(h), 0, -5, 50.0, 0.03, 0) == -1  # bad capacity
    assert L.tg_check_configs(None, None, None, 0, None, None, None) == -1


def test_sass_is_sm100a(built_lib):
    """The shipped cubin targets sm_100a only (no PTX for other archs, no multi-backend)."""
    import shutil
    import subprocess

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", built_lib[0]], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_types_and_factories():
    import tygrit_b200 as tb
    from tygrit_b200 import planner as P
    from tygrit_b200.types import base_xyt, make_trajectory

    assert base_xyt(tb.SE2Pose(1, 2, 3)) == (1.0, 2.0, 3.0) and base_xyt([4, 5, 6]) == (4.0, 5.0, 6.0)
    tr = make_trajectory([[0] * 8, [1] * 8], [[0, 0, 0], [1, 2, 3]])
    assert tr.arm_path[1].shape == (8,) and tr.base_configs[1].theta == 3.0
    with pytest.raises(ValueError, match="Unknown robot"):
        P.create_planner("nonexistent", "b200")
    with pytest.raises(ValueError, match="Unknown Fetch planner"):
        P.create_planner("fetch", "nonexistent")
    with pytest.raises(TypeError, match="Expected B200PlannerConfig"):
        P.create_planner("fetch", "b200", config=tb.PlannerConfig())
    assert tb.PlannerFailure.COLLISION_AT_START.value == "collision_at_start"
    assert tb.B200PlannerConfig().point_radius == 0.03 and tb.B200PlannerConfig().interpolation_steps == 16


def test_halton_and_interpolation():
    from tygrit_b200.planner import B200Planner, _halton

    h = _halton(4, 2, 0)
    np.testing.assert_allclose(h[:, 0], [0.5, 0.25, 0.75, 0.125])
    np.testing.assert_allclose(h[:3, 1], [1 / 3, 2 / 3, 1 / 9])
    p = np.array([[0.0] * 11, [1.0] * 11])
    out = B200Planner._interpolate_steps(p, 4)
    assert out.shape == (5, 11) and np.allclose(out[2], 0.5)
    out = B200Planner._interpolate_density(p, 0.5)
    assert np.all(np.linalg.norm(np.diff(out, axis=0), axis=1) <= 0.5 + 1e-9)


def test_synth_is_deterministic():
    from tygrit_b200 import synth

    a = synth.room_cloud(5000)
    b = synth.room_cloud(5000)
    assert a.dtype == np.float32 and a.shape == (5000, 3) and np.array_equal(a, b)
    assert a[:, 2].min() > synth.GROUND_Z
    c = synth.room_cloud(5000, perturb_step=3)
    assert not np.array_equal(a, c)
    arm, base = synth.random_configs(100)
    assert arm.shape == (100, 8) and base.shape == (100, 3)
    st = synth.random_states(5)
    assert st.shape == (5, 13)
    t = synth.tabletop_cloud(1000)
    assert t.shape == (1000, 3) and abs(np.median(t[:500, 2]) - 0.72) < 0.01


def test_product_does_not_import_oracle():
    """The product path must never route through the oracle (it would void every parity claim)."""
    pkg = os.path.join(ROOT, "tygrit_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "libtg_oracle" not in txt.replace(
                    'os.path.join(ROOT, "oracle", "libtg_oracle.so")', ""
                ), f
