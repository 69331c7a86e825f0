"""CPU oracle for the TyGrit collision-checking hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``tygrit_b200/`` may import this
package: it is the checker, never the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs call into it.

What it restates (every function cites the reference file:line it follows,
paths relative to the upstream TyGrit tree):

* ``fetch_model``  – the Fetch kinematic constants and the 111-sphere table
  (``TyGrit/kinematics/fetch/constants.py:17-31,74-79,84-225``), by value.
* ``fk``           – fp64 link frames and world sphere centres
  (``TyGrit/kinematics/fetch/fk_numpy.py:32-157``,
  ``TyGrit/utils/transforms.py:116-127``).
* ``collide``      – exact sphere-vs-cloud clearance / verdict, edge sampling
  and edge verdicts (the arithmetic lives in third-party VAMP,
  ``AdaCompNUS/vamp@ab0bc393107ebcf445a2e8df52c1aec7835ace4e``, which is NOT
  in the reference tree; anchored on the call sites
  ``TyGrit/planning/fetch/vamp_preview.py:159-163,183-236,306-338``).
* ``ingest``       – robot self-filter and depth→world cloud
  (``TyGrit/kinematics/robot_filter.py:13-64``, ``TyGrit/utils/depth.py:13-84``,
  ``TyGrit/utils/pointcloud.py:13-90``).
* ``cpu_checker.c``– the same verdict arithmetic in C (OpenMP), used for
  full-size parity runs and as the timed CPU baseline.

Pinning status
--------------
* FK frames, sphere centres, camera pose, self-filter, depth ingest:
  PINNED — checked against the reference's own Python (imported from
  ``/root/reference`` by ``tools/make_golden.py``; vectors committed under
  ``tests/golden/``) and against the reference tests' known-answer values
  (SAPIEN fixtures ``test/test_kinematics.py:208-577``).
* Collision verdicts / edge verdicts / clearance vs VAMP: **parity unpinned**.
  VAMP's source is absent, not installed and cannot be fetched; no reference
  test pins a verdict against a cloud (``test/test_planner.py:72-78`` only
  checks an empty environment).  The contract this oracle *defines* is spelled
  out in ``collide.py`` and DESIGN.md so that a later run with VAMP available
  can diff it.
"""
