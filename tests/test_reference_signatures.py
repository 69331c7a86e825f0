"""CPU: the drop-in boundary checked against the REFERENCE's own protocols and classes, not against itself.

Imports the unmodified reference from ``/root/reference`` (present in the build container only; skipped elsewhere)
and compares, by ``inspect.signature``, method names, parameter names, parameter kinds and defaults of

* ``MotionPlanner`` (``TyGrit/planning/motion_planner.py:15-64``) and every public method of
  ``VampPreviewPlanner`` (``planning/fetch/vamp_preview.py:72-322``)          vs ``B200Planner``
* ``Scene`` (``belief_state/scene.py:17-48``) and ``PointCloudScene``            vs ``B200Scene``
* ``BatchFKSolver`` (``kinematics/fk.py:84-106``)                                vs ``B200BatchFK``
* ``filter_robot_points``, ``check_self_occlusion``, the factories              vs their twins
* ``VampPlannerConfig`` fields and defaults (``planning/config.py:9-47``)        vs ``B200PlannerConfig``
* the value types: with TyGrit importable, ``tygrit_b200.types`` re-exports TyGrit's own classes (enum identity).
"""

from __future__ import annotations

import dataclasses
import importlib
import inspect
import os
import sys

import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "TyGrit")), reason="reference tree not present on this box")


@pytest.fixture(scope="module")
def ref():
    sys.path.insert(0, REF)
    try:
        yield importlib.import_module("TyGrit")
    finally:
        sys.path.remove(REF)


def _params(fn):
    return [(p.name, p.kind, p.default) for p in inspect.signature(fn).parameters.values()]


def _assert_compatible(theirs, ours, what):
    """Every call that is valid for the reference's function is valid for ours and means the same: same parameter names,
    order and kinds; a default of the reference is kept; ours may only ADD defaults (or trailing defaulted parameters)."""
    a, b = _params(theirs), _params(ours)
    assert len(b) >= len(a), f"{what}: {inspect.signature(ours)} has fewer parameters than {inspect.signature(theirs)}"
    for (n1, k1, d1), (n2, k2, d2) in zip(a, b):
        assert (n1, k1) == (n2, k2), f"{what}: parameter {n2!r} where the reference has {n1!r}"
        assert d1 is inspect.Parameter.empty or d1 == d2, f"{what}: default of {n1!r} is {d2!r}, reference {d1!r}"
    for n2, k2, d2 in b[len(a):]:
        assert d2 is not inspect.Parameter.empty or k2 in (inspect.Parameter.VAR_KEYWORD, inspect.Parameter.VAR_POSITIONAL), (
            f"{what}: extra parameter {n2!r} without a default")


def _public_methods(cls):
    return {n: f for n, f in inspect.getmembers(cls, lambda o: inspect.isfunction(o) or isinstance(o, property)) if not n.startswith("_")}


def _assert_same_surface(ref_cls, ours, names=None):
    want = _public_methods(ref_cls)
    have = _public_methods(ours)
    for name in (names or want):
        assert name in have, f"{ours.__name__} lacks {ref_cls.__name__}.{name}"
        if isinstance(want[name], property):
            assert isinstance(have[name], property), name
            continue
        _assert_compatible(want[name], have[name], f"{ours.__name__}.{name}")


def test_planner_surface_matches_reference(ref):
    from TyGrit.planning.fetch.vamp_preview import VampPreviewPlanner
    from TyGrit.planning.motion_planner import MotionPlanner

    from tygrit_b200.planner import B200Planner

    _assert_same_surface(MotionPlanner, B200Planner)
    _assert_same_surface(VampPreviewPlanner, B200Planner)            # incl. the concrete-only helpers
    _assert_compatible(VampPreviewPlanner.__init__, B200Planner.__init__, "B200Planner.__init__")


def test_scene_and_fk_surfaces_match_reference(ref):
    from TyGrit.belief_state.pointcloud_scene import PointCloudScene
    from TyGrit.belief_state.scene import Scene
    from TyGrit.kinematics.fk import BatchFKSolver

    from tygrit_b200.fk import B200BatchFK
    from tygrit_b200.scene import B200Scene

    _assert_same_surface(Scene, B200Scene)
    _assert_same_surface(PointCloudScene, B200Scene)
    _assert_compatible(PointCloudScene.__init__, B200Scene.__init__, "B200Scene.__init__")   # ours adds keyword-only extras
    _assert_same_surface(BatchFKSolver, B200BatchFK)


def test_functions_and_factories_match_reference(ref):
    from TyGrit.checker.collision import create_collision_check as r_ccc
    from TyGrit.checker.occlusion import check_self_occlusion as r_occ
    from TyGrit.kinematics.fk import create_fk_solver as r_fk
    from TyGrit.kinematics.robot_filter import filter_robot_points as r_filter
    from TyGrit.planning.fetch import create_fetch_planner as r_cfp
    from TyGrit.planning.planner import create_planner as r_cp

    from tygrit_b200 import fk, occlusion, planner, robot_filter

    for theirs, ours in ((r_filter, robot_filter.filter_robot_points), (r_occ, occlusion.check_self_occlusion),
                         (r_cp, planner.create_planner), (r_cfp, planner.create_fetch_planner),
                         (r_ccc, planner.create_collision_check), (r_fk, fk.create_fk_solver)):
        _assert_compatible(theirs, ours, f"{ours.__module__}.{ours.__name__}")
    # error behaviour of the factories (planning/fetch/__init__.py:25-32, planning/planner.py:36-39)
    with pytest.raises(ValueError):
        planner.create_planner("pr2", "b200")
    with pytest.raises(ValueError):
        planner.create_fetch_planner("nope")
    with pytest.raises(TypeError):
        planner.create_fetch_planner("b200", object())


def test_config_fields_and_defaults_match_reference(ref):
    from TyGrit.planning.config import PlannerConfig as RPC, VampPlannerConfig

    from tygrit_b200.config import B200PlannerConfig, PlannerConfig

    ours = {f.name: f.default for f in dataclasses.fields(B200PlannerConfig)}
    for f in dataclasses.fields(VampPlannerConfig):
        assert f.name in ours, f"B200PlannerConfig lacks VampPlannerConfig.{f.name}"
        assert ours[f.name] == f.default, f"default of {f.name}: {ours[f.name]!r} != {f.default!r}"
    assert {f.name: f.default for f in dataclasses.fields(PlannerConfig)} == {f.name: f.default for f in dataclasses.fields(RPC)}


def test_value_types_are_tygrits_own_when_importable(ref):
    import tygrit_b200.types as t
    from TyGrit.types.failures import PlannerFailure
    from TyGrit.types.geometry import SE2Pose
    from TyGrit.types.planning import Trajectory
    from TyGrit.types.results import PlanResult
    from TyGrit.types.robot import WholeBodyConfig

    ext = t.tygrit_types()                      # TyGrit is importable inside this fixture
    assert ext is not None
    assert ext["PlannerFailure"] is PlannerFailure and ext["PlanResult"] is PlanResult and ext["Trajectory"] is Trajectory
    assert ext["SE2Pose"] is SE2Pose and ext["WholeBodyConfig"] is WholeBodyConfig
    tr = t.make_trajectory([[0.0] * 8], [[1.0, 2.0, 0.5]])
    assert isinstance(tr, t.Trajectory) and isinstance(tr.base_configs[0], t.SE2Pose)
    if t.USING_TYGRIT_TYPES:                    # tygrit_b200 was first imported with TyGrit on the path
        assert t.PlanResult is PlanResult and t.PlannerFailure is PlannerFailure


def test_structural_twins_have_the_reference_fields(ref):
    """Standalone, the twins are used: same field names and order, same enum values."""
    import tygrit_b200.types as t

    pairs = [("TyGrit.types.geometry", "SE2Pose"), ("TyGrit.types.robot", "WholeBodyConfig"), ("TyGrit.types.robot", "RobotState"),
             ("TyGrit.types.sensor", "SensorSnapshot"), ("TyGrit.types.planning", "Trajectory"), ("TyGrit.types.results", "PlanResult")]
    for mod, name in pairs:
        theirs = getattr(importlib.import_module(mod), name)
        ours = t._TWINS[name]
        assert [f.name for f in dataclasses.fields(theirs)] == [f.name for f in dataclasses.fields(ours)], name
    from TyGrit.types.failures import PlannerFailure

    assert {m.name: m.value for m in PlannerFailure} == {m.name: m.value for m in t._TWINS["PlannerFailure"]}
