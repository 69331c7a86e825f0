"""GPU: parity of the CUDA path (through the C ABI) with the oracle.

Bar: verdicts / masks bit-exact for every configuration, edge or point whose fp64 clearance
satisfies |clearance| > 1e-5 m; FK sphere centres within 1e-6 m (fp32) — BASELINE.json north_star.
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import cchecker, collide, fk as ofk, ingest as oin
from tygrit_b200 import engine, synth
from tygrit_b200.engine import CollisionWorld

pytestmark = pytest.mark.gpu
BAND = 1e-5
FK_TOL = 1e-6


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def tabletop():
    cloud = synth.tabletop_cloud(50_000)
    w = CollisionWorld(max_points=1 << 20)
    w.build(_cuda(cloud))
    return cloud, w, cchecker.CCloud(cloud)


@pytest.fixture(scope="module")
def room():
    cloud = synth.room_cloud(200_000)
    w = CollisionWorld(max_points=1 << 20)
    w.build(_cuda(cloud))
    return cloud, w, cchecker.CCloud(cloud)


# ------------------------------------------------------------------ FK
def test_fk_frames_vs_reference_golden(golden):
    g = golden("fk.npz")
    fr = engine.fk_frames(_cuda(g["q10"].astype(np.float32))).cpu().numpy()
    assert fr.shape == (len(g["q10"]), 15, 4, 4)
    ref = ofk.link_frames_4x4(g["q10"].astype(np.float32))     # same f32-rounded inputs, fp64 maths
    assert np.abs(fr - ref).max() <= FK_TOL
    assert np.abs(fr - g["frames"]).max() <= 1e-5              # reference's own torch-vs-numpy tolerance


def test_fk_sphere_centres_within_1e6():
    arm, base = synth.random_configs(100_000)
    c = engine.fk_spheres(_cuda(arm), _cuda(base)).cpu().numpy()
    ref = cchecker.centres_f64(arm, base)
    err = np.abs(c - ref).max()
    assert err <= FK_TOL, err
    st = synth.random_states(4096)
    c = engine.fk_spheres(_cuda(st[:, 3:11]), _cuda(st[:, :3]), _cuda(st[:, 11:13])).cpu().numpy()
    ref = cchecker.centres_f64(st[:, 3:11], st[:, :3], st[:, 11:13])
    assert np.abs(c - ref).max() <= FK_TOL


def test_batch_fk_solver_protocol(golden):
    from tygrit_b200.fk import create_fk_solver

    g = golden("fk.npz")
    solver = create_fk_solver("fetch", "fk_batch")
    assert solver.base_frame == "base_link"
    out = solver.solve(_cuda(g["q10"].astype(np.float32)))
    assert set(out) == set(g["link_names"]) and out["gripper_link"].shape == (len(g["q10"]), 4, 4)
    np.testing.assert_allclose(out["gripper_link"][0, :3, 3].cpu().numpy(), [1.1281, 0, 0.78601], atol=1e-4)
    with pytest.raises(RuntimeError):
        solver.solve(torch.zeros(2, 10))


# ------------------------------------------------------------------ config verdicts
def _assert_verdicts(valid, cl, what=""):
    valid = np.asarray(valid).astype(bool)
    decisive = np.abs(cl) > BAND
    bad = np.flatnonzero((valid != (cl > 0)) & decisive)
    assert bad.size == 0, f"{what}: {bad.size} verdict mismatches outside the band, e.g. idx {bad[:5]} clearance {cl[bad[:5]]}"


def test_c1_tabletop_4096_configs(tabletop):
    """BASELINE config[0]: 4096 random configs vs the 50k-point tabletop cloud."""
    cloud, w, cc = tabletop
    arm, base = synth.random_configs(4096, base_box=(-1, 1, -1, 1))
    cl = collide.CloudOracle(cloud).config_clearance(arm, base)         # numpy/cKDTree oracle
    valid, clr = w.check_configs(_cuda(arm), _cuda(base), clearance=True)
    _assert_verdicts(valid.cpu().numpy(), cl, "C1")
    assert 0.3 < valid.float().mean().item() < 0.8
    got = clr.cpu().numpy()
    np.testing.assert_allclose(got, np.minimum(cl, 0.25), atol=3e-6)
    c = w.counters()
    assert c["narrow_items"] > 0 and c["definite_collisions"] > 0     # both phases were exercised


def test_room_configs_vs_c_oracle(room):
    cloud, w, cc = room
    arm, base = synth.random_configs(50_000, seed=7)
    cl = cc.clearance(arm, base)
    valid = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    _assert_verdicts(valid, cl, "room")
    assert 0.05 < valid.mean() < 0.6


def test_queue_overflow_resolves_inline_with_same_verdicts(room):
    """A launch pair takes up to 64 Mi configurations against an 8 Mi-item queue; items that do not fit are resolved
    inline by their own lane.  Shrink the queue (test hook) so that most items overflow: verdicts and edge verdicts must
    not change, and the overflow counter must say the path ran."""
    cloud, w, cc = room
    arm, base = synth.random_configs(60_000, seed=17)
    a0, b0, a1, b1 = synth.random_edges(2_000, seed=18)
    w.counters()
    want = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    want_e = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1)).cpu().numpy()
    assert w.counters()["overflow_inline"] == 0
    try:
        w.set_queue_limit(64)
        got = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
        c = w.counters()
        got_e = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1)).cpu().numpy()
        assert c["overflow_inline"] > 100
        assert w.counters()["overflow_inline"] > 100
    finally:
        w.set_queue_limit(0)
    assert np.array_equal(got, want) and np.array_equal(got_e, want_e)
    _assert_verdicts(got, cc.clearance(arm, base), "overflow")


def test_ragged_and_degenerate_batches(room):
    cloud, w, cc = room
    arm, base = synth.random_configs(1000, seed=3)
    full = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    for n in (0, 1, 31, 33, 127, 129, 999):
        part = w.check_configs(_cuda(arm[:n]), _cuda(base[:n])).cpu().numpy()
        assert np.array_equal(part, full[:n])
    # far outside the grid: nothing to hit
    far = base.copy()
    far[:, :2] += 500.0
    assert w.check_configs(_cuda(arm), _cuda(far)).all()
    # NaN configuration must not crash (verdict unspecified)
    bad = arm.copy()
    bad[0, 3] = np.nan
    w.check_configs(_cuda(bad), _cuda(base))
    torch.cuda.synchronize()


def test_batch_sizes_at_the_launch_shape_boundaries(tabletop):
    """Batch sizes on both sides of every boundary of the configuration kernel's launch shape: one sub-block tile (512), one
    three-tile stage-A1 step (1536), the switch to the coarse-grid variant (2^18) and one full step of every sub-block of
    every SM (3 x 512 x 2 x SM count); every configuration against the fp64 oracle, and a prefix of a large batch must get
    the verdicts it gets alone (the ring carries survivors across steps)."""
    cloud, w, cc = tabletop
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    full = 3 * 512 * 2 * sms
    arm, base = synth.random_configs(full + 1537, seed=77, base_box=(-1.2, 1.2, -1.2, 1.2))
    cl = cc.clearance(arm, base, cap=1e-3)
    a, b = _cuda(arm), _cuda(base)
    ref_big = w.check_configs(a, b).cpu().numpy()
    _assert_verdicts(ref_big, cl, "largest batch")
    for n in (1, 511, 512, 513, 1535, 1536, 1537, 3071, 3073, (1 << 18) - 1, 1 << 18, (1 << 18) + 1, full - 1, full, full + 1):
        got = w.check_configs(a[:n], b[:n]).cpu().numpy()
        _assert_verdicts(got, cl[:n], f"batch of {n}")
        dec = np.abs(cl[:n]) > BAND
        assert np.array_equal(got[dec], ref_big[:n][dec]), f"batch of {n}: verdicts differ from the same configurations in the large batch"


def test_empty_single_and_nonfinite_clouds():
    w = CollisionWorld(max_points=4096)
    arm, base = synth.random_configs(500, seed=5)
    w.build(np.empty((0, 3), np.float32))
    assert w.check_configs(_cuda(arm), _cuda(base)).all()           # empty world: everything is free
    # tuck configuration at the origin is valid in an empty environment (test/test_planner.py:72-78)
    tuck = np.array([[0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]], np.float32)
    assert w.check_configs_host(tuck, np.zeros((1, 3), np.float32))[0] == 1
    one = np.array([[0.9, 0.0, 0.6]], np.float32)
    w.build(one)
    cl = collide.CloudOracle(one).config_clearance(arm, base * np.array([0.3, 0.3, 1], np.float32))
    v = w.check_configs(_cuda(arm), _cuda(base * np.array([0.3, 0.3, 1], np.float32))).cpu().numpy()
    _assert_verdicts(v, cl, "single point")
    assert (cl <= 0).any()
    dirty = np.vstack([one, [[np.nan, 0, 0]], [[np.inf, 1, 1]]]).astype(np.float32)
    w.build(dirty)
    v2 = w.check_configs(_cuda(arm), _cuda(base * np.array([0.3, 0.3, 1], np.float32))).cpu().numpy()
    assert np.array_equal(v, v2) and w.stats()["n_dropped"] == 2
    w.close()


def test_obstacles_and_attachment(tabletop):
    cloud, _, _ = tabletop
    w = CollisionWorld(max_points=1 << 17)
    w.build(cloud)
    arm, base = synth.random_configs(3000, seed=11, base_box=(-1, 1, -1, 1))
    obs = np.array([[0.6, 0.3, 1.0, 0.12], [0.2, -0.5, 0.5, 0.2]], np.float32)
    att = {"spheres": np.array([[0.1, 0.0, 0.0, 0.04], [0.18, 0.02, 0.0, 0.03]]), "pos": (0.02, 0.0, 0.01),
           "quat": (0.0, 0.3826834, 0.0, 0.9238795)}
    w.set_obstacles(obs)
    w.set_attachment(att["spheres"], att["pos"], att["quat"])
    orc = collide.CloudOracle(cloud, obstacles=obs)
    cl = orc.config_clearance(arm, base, attachment=att)
    cl_plain = collide.CloudOracle(cloud).config_clearance(arm, base)
    v = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    _assert_verdicts(v, cl, "extras")
    assert ((cl <= 0) & (cl_plain > 0)).sum() > 10                  # the extras changed some verdicts
    w.set_obstacles(np.empty((0, 4), np.float32))
    w.set_attachment(np.empty((0, 4), np.float32))
    _assert_verdicts(w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), cl_plain, "extras cleared")
    w.close()


def test_clearance_with_obstacles_attachment_and_self_term(tabletop):
    """The clearance output when obstacle / attachment spheres are set (the lane-per-configuration kernel serves these worlds)
    and with the self-collision term: within 3e-6 m of min(cloud + extras clearance, self clearance, 0.25) in fp64."""
    cloud, _, _ = tabletop
    w = CollisionWorld(max_points=1 << 17)
    w.build(cloud)
    arm, base = synth.random_configs(3000, seed=21, base_box=(-1.2, 1.2, -1.2, 1.2))
    obs = np.array([[0.6, 0.3, 1.0, 0.12], [0.2, -0.5, 0.5, 0.2], [-0.9, 0.9, 0.3, 0.05]], np.float32)
    att = {"spheres": np.array([[0.1, 0.0, 0.0, 0.04], [0.18, 0.02, 0.0, 0.03]]), "pos": (0.02, 0.0, 0.01),
           "quat": (0.0, 0.3826834, 0.0, 0.9238795)}
    w.set_obstacles(obs)
    w.set_attachment(att["spheres"], att["pos"], att["quat"])
    cl = collide.CloudOracle(cloud, obstacles=obs).config_clearance(arm, base, attachment=att)
    for self_on in (False, True):
        w.set_self_collision(self_on)
        ref = np.minimum(cl, cchecker.self_clearance(arm)) if self_on else cl
        v, c = w.check_configs(_cuda(arm), _cuda(base), clearance=True)
        _assert_verdicts(v.cpu().numpy(), ref, f"extras, self={self_on}")
        np.testing.assert_allclose(c.cpu().numpy(), np.minimum(ref, 0.25), atol=3e-6, rtol=0)
    w.close()


def test_point_radius_and_coarse_grid(tabletop):
    cloud, _, cc = tabletop
    arm, base = synth.random_configs(4000, seed=13, base_box=(-1, 1, -1, 1))
    for vpm, r_pt in ((25.0, 0.03), (50.0, 0.0), (80.0, 0.05)):
        w = CollisionWorld(max_points=1 << 16, voxels_per_metre=vpm, point_radius=r_pt)
        w.build(cloud)
        cl = cchecker.CCloud(cloud, point_radius=r_pt).clearance(arm, base)
        _assert_verdicts(w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), cl, f"vpm={vpm} r={r_pt}")
        w.close()
    # a world too big for the requested resolution is coarsened, never truncated
    w = CollisionWorld(max_points=1 << 16, voxels_per_metre=100.0, max_voxels=1 << 20)
    w.build(cloud)
    assert w.stats()["voxels_per_metre"] < 100.0
    _assert_verdicts(w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), cc.clearance(arm, base), "coarsened")
    w.close()


def test_cell_shape_does_not_change_verdicts_or_clearances(tabletop):
    """The narrow-phase cell grid (isotropic, the anisotropic default, extreme aspect ratios, cells larger than the
    largest sphere) only decides which candidate points are read: verdicts and clearances must be identical."""
    cloud, w0, cc = tabletop
    arm, base = synth.random_configs(20_000, seed=17, base_box=(-1, 1, -1, 1))
    cl = cc.clearance(arm, base)
    v0, c0 = w0.check_configs(_cuda(arm), _cuda(base), clearance=True)
    _assert_verdicts(v0.cpu().numpy(), cl, "default cell shape")
    a0, b0 = synth.random_configs(3000, seed=18, base_box=(-1, 1, -1, 1))
    a1, _ = synth.random_configs(3000, seed=19)
    e0 = w0.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b0))
    for shape in ((20.0, 20.0), (160.0, 4.0), (3.0, 60.0), (2.0, 2.0), (200.0, 100.0)):
        w = CollisionWorld(max_points=1 << 16)
        w.set_cell_shape(*shape)
        w.build(cloud)
        v, c = w.check_configs(_cuda(arm), _cuda(base), clearance=True)
        assert torch.equal(v, v0), shape
        assert torch.equal(c, c0), shape
        assert torch.equal(w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b0)), e0), shape
        assert w.counters()["narrow_items"] > 0
        w.close()


def test_host_api_matches_device_api_and_chunks(room):
    cloud, w, cc = room
    arm, base = synth.random_configs(1_300_000, seed=21)              # > one 1M pipeline chunk
    dev = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    host = w.check_configs_host(arm, base)
    assert np.array_equal(dev, host)
    sub = np.random.default_rng(0).choice(len(arm), 20_000, replace=False)
    _assert_verdicts(host[sub], cc.clearance(arm[sub], base[sub]), "host api")
    torch.cuda.synchronize()
    again = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    assert np.array_equal(dev, again)                                  # deterministic


def test_host_clearances_right_after_a_build_across_pipeline_chunks():
    """The half-resolution grid of the clearance query is built lazily by the FIRST clearance call after a build; the host
    pipeline alternates between two streams, so the second chunk must not read it before the first chunk's stream has
    finished building it.  Fresh world, 2.3 M configurations (three chunks), clearances vs the device path and the oracle."""
    cloud = synth.room_cloud(150_000, seed=5)
    w = CollisionWorld(max_points=1 << 18, voxels_per_metre=64.0)
    w.build(_cuda(cloud))
    arm, base = synth.random_configs(2_300_000, seed=23)
    hv, hc = w.check_configs_host(arm, base, clearance=True)           # first clearance call on this structure
    dv, dc = w.check_configs(_cuda(arm), _cuda(base), clearance=True)
    assert np.array_equal(hv, dv.cpu().numpy())
    assert np.array_equal(hc, dc.cpu().numpy())
    sub = np.concatenate([np.arange(1_040_000, 1_060_000), np.arange(2_090_000, 2_110_000)])   # around the chunk boundaries
    ref = cchecker.CCloud(cloud).clearance(arm[sub], base[sub], cap=0.25)
    np.testing.assert_allclose(hc[sub], ref, atol=3e-6, rtol=0)
    w.close()


def _read_structure(w):
    """Read the distance grid and the coarse far-field grid back from the structure blob (tg_env_grid_info gives the layout)."""
    gi = w.grid_info()
    blob = w.blob_view(w.used_bytes()).cpu().numpy()
    nx, ny, nz = int(gi["nx"]), int(gi["ny"]), int(gi["nz"])
    gnx, gny, gnz = int(gi["gnx"]), int(gi["gny"]), int(gi["gnz"])
    off_df, off_cg = int(gi["off_df"]), int(gi["off_cg"])
    df = blob[off_df:off_df + nx * ny * nz].reshape(nz, ny, nx)
    cg = blob[off_cg:off_cg + gnx * gny * gnz].reshape(gnz, gny, gnx)
    return dict(vpm=np.float32(gi["voxels_per_metre"]), bias=np.array([gi["bias_x"], gi["bias_y"], gi["bias_z"]], np.float32), n=(nx, ny, nz),
                q=gi["q"], slack=gi["slack"], W=int(gi["window"]), df=df, cg=cg, cg_r=int(gi["cg_r"]), cg_q=gi["cg_q"])


@pytest.mark.parametrize("vpm", [50.0, 80.0])
def test_distance_grid_codes_bit_exact_vs_scipy_edt(vpm):
    """K1's truncated EDT (bitmask X pass, DPX add-min Y/Z passes) against scipy's exact transform:
    every voxel code must be floor(sqrt(d2) * 255 / W) of the exact squared voxel-centre distance."""
    from scipy import ndimage

    rng = np.random.default_rng(5)
    pts = np.concatenate([
        synth.tabletop_cloud(6000),
        rng.uniform([-0.4, -0.9, 0.1], [1.9, 0.9, 1.7], (300, 3)).astype(np.float32),   # isolated points in free space
    ]).astype(np.float32)
    w = CollisionWorld(max_points=1 << 16, voxels_per_metre=vpm)
    w.build(_cuda(pts))
    torch.cuda.synchronize()
    st = _read_structure(w)
    nx, ny, nz = st["n"]
    W = st["W"]
    assert W <= 31
    # voxel of a point: round-to-nearest of fma(x, vpm, bias) - 2^23 (same as k_count); ties are measure zero
    f = (pts.astype(np.float64) * np.float64(st["vpm"]) + st["bias"].astype(np.float64)).astype(np.float32)
    idx = (f.astype(np.float64) - 8388608.0).astype(np.int64)
    assert (idx >= 0).all() and (idx < np.array([nx, ny, nz])).all()
    occ = np.zeros((nz, ny, nx), dtype=bool)
    occ[idx[:, 2], idx[:, 1], idx[:, 0]] = True
    d2 = np.rint(ndimage.distance_transform_edt(~occ) ** 2).astype(np.int64)
    code = np.floor(np.sqrt(d2.astype(np.float32)) * (np.float32(255.0) / np.float32(W))).astype(np.int64)
    want = np.where(d2 < W * W, np.minimum(code, 254), 255).astype(np.uint8)
    got = st["df"]
    bad = np.argwhere(got != want)
    assert bad.shape[0] == 0, f"{bad.shape[0]} voxel codes differ, first {bad[:3]}: got {got[tuple(bad[0])]} want {want[tuple(bad[0])]}"
    # coarse far-field grid: every code must be a LOWER bound (in units of cg_q) on the distance from ANY point of its
    # block of cg_r^3 voxels to the cloud, and not uselessly loose.  Exact reference: distance of the voxel centres to the
    # points' voxel centres (scipy) minus the two half-diagonals a point and a query can sit off their voxel centres.
    r, cq, cg = st["cg_r"], st["cg_q"], st["cg"]
    assert cg.shape == (-(-nz // r), -(-ny // r), -(-nx // r)) and cg.size <= 180224
    d = ndimage.distance_transform_edt(~occ) / np.float64(st["vpm"])
    pad = [(0, cg.shape[a] * r - d.shape[a]) for a in range(3)]
    dmin = np.pad(d, pad, constant_values=np.inf).reshape(cg.shape[0], r, cg.shape[1], r, cg.shape[2], r).min(axis=(1, 3, 5))
    true_lo = dmin - np.sqrt(3.0) / np.float64(st["vpm"])          # any point of a voxel vs any point of an occupied voxel
    claimed = cg.astype(np.float64) * cq
    assert (claimed <= np.maximum(true_lo, 0.0) + 1e-6).all(), "a coarse code claims more clearance than there is"
    far = dmin > 3.5 * r / np.float64(st["vpm"])
    assert far.any() and (claimed[far] >= np.minimum(dmin[far], 1.0) - 2.0 * np.sqrt(3.0) * r / np.float64(st["vpm"]) - 2 * cq).all()
    w.close()


def test_structure_export_import(room):
    cloud, w, cc = room
    arm, base = synth.random_configs(20_000, seed=31)
    ref = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    n = w.used_bytes()
    w2 = CollisionWorld(max_points=1 << 20)
    w2.blob_view(n).copy_(w.blob_view(n))                              # what an NCCL broadcast would do
    torch.cuda.synchronize()
    w2.adopt_blob(n)
    assert np.array_equal(w2.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), ref)
    w3 = CollisionWorld(max_points=1 << 20)
    w3.import_blob(w.blob_view(n).clone())
    assert np.array_equal(w3.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), ref)
    w2.close()
    w3.close()


# ------------------------------------------------------------------ edges
def test_c2_edges_vs_oracle(room):
    """BASELINE config[1] at oracle-sized scale: random Fetch edges at VAMP default resolution."""
    cloud, w, cc = room
    a0, b0, a1, b1 = synth.random_edges(3000, seed=42)
    cl, ns = cc.edge_clearance(a0, b0, a1, b1)
    valid, nchk = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1), count=True)
    assert int(nchk.item()) == int(ns.sum())
    _assert_verdicts(valid.cpu().numpy(), cl, "C2 edges")
    # short RRT-like edges + whole-body edges (base moves too)
    a0, b0, a1, b1 = synth.random_edges(3000, seed=43, rrt_like=True)
    b1 = b1 + np.random.default_rng(1).normal(0, 0.1, b1.shape).astype(np.float32)
    cl, ns = cc.edge_clearance(a0, b0, a1, b1)
    valid, nchk = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1), count=True)
    assert int(nchk.item()) == int(ns.sum())
    _assert_verdicts(valid.cpu().numpy(), cl, "short edges")
    assert 0 < valid.sum() < len(valid)
    hv, hn = w.check_edges_host(a0, b0, a1, b1)
    assert np.array_equal(hv, valid.cpu().numpy()) and hn == int(ns.sum())


def test_edge_properties(room):
    cloud, w, cc = room
    arm, base = synth.random_configs(5000, seed=51)
    cfg_ok = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    # a zero-length edge samples its end point rake times: same verdict as the configuration
    e_ok = w.check_edges(_cuda(arm), _cuda(base), _cuda(arm), _cuda(base)).cpu().numpy()
    assert np.array_equal(cfg_ok, e_ok)
    # a valid edge implies a valid end point (end included, start excluded)
    a0, b0, a1, b1 = synth.random_edges(5000, seed=52, rrt_like=True)
    ev = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1)).cpu().numpy().astype(bool)
    end_ok = w.check_configs(_cuda(a1), _cuda(b1)).cpu().numpy().astype(bool)
    assert not (ev & ~end_ok).any()


def test_full_size_chunk_independence(room):
    """BASELINE-size batches cross the library's launch boundaries (2^19 edges, 64 Mi configurations per launch pair):
    the verdict of a unit must not depend on where in the batch it sits.  Size-independent property: a permuted batch
    gives the permuted verdicts, and a slice re-checked on its own gives the same bytes."""
    cloud, w, cc = room
    n_e = 1_200_000                                                    # 3 edge launches (2^19 edges each)
    a0, b0, a1, b1 = synth.random_edges(n_e, seed=61, rrt_like=True)
    ev, nchk = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1), count=True)
    ev = ev.cpu().numpy()
    perm = np.random.default_rng(62).permutation(n_e)
    ev_p, nchk_p = w.check_edges(_cuda(a0[perm]), _cuda(b0[perm]), _cuda(a1[perm]), _cuda(b1[perm]), count=True)
    assert np.array_equal(ev_p.cpu().numpy(), ev[perm]) and int(nchk.item()) == int(nchk_p.item())
    sl = slice((1 << 19) - 1000, (1 << 19) + 1000)                     # straddles the first launch boundary
    ev_s = w.check_edges(_cuda(a0[sl]), _cuda(b0[sl]), _cuda(a1[sl]), _cuda(b1[sl])).cpu().numpy()
    assert np.array_equal(ev_s, ev[sl])
    cl, _ = cc.edge_clearance(a0[sl], b0[sl], a1[sl], b1[sl])
    _assert_verdicts(ev_s, cl, "edges across a launch boundary")
    # configurations: 72 Mi = two launch pairs (64 Mi + 8 Mi); generated on the device, checked by slices + the C oracle
    B = 72 * (1 << 20)
    g = torch.Generator(device="cuda")
    g.manual_seed(63)
    lo = torch.tensor(synth.JOINT_LOWER, dtype=torch.float32, device="cuda")
    hi = torch.tensor(synth.JOINT_UPPER, dtype=torch.float32, device="cuda")
    arm = torch.rand(B, 8, generator=g, device="cuda") * (hi - lo) + lo
    blo = torch.tensor([-3.5, -2.5, -np.pi], dtype=torch.float32, device="cuda")
    bhi = torch.tensor([3.5, 2.5, np.pi], dtype=torch.float32, device="cuda")
    base = torch.rand(B, 3, generator=g, device="cuda") * (bhi - blo) + blo
    valid = w.check_configs(arm, base)
    assert w.counters()["overflow_inline"] == 0
    for start in (0, (32 << 20) - 1000, (64 << 20) - 1000, B - 2000):  # head, middle, the launch boundary, tail
        s2 = slice(start, start + 2000)
        again = w.check_configs(arm[s2].contiguous(), base[s2].contiguous())
        assert torch.equal(again, valid[s2])
        cl = cc.clearance(arm[s2].cpu().numpy(), base[s2].cpu().numpy())
        _assert_verdicts(valid[s2].cpu().numpy(), cl, "configs at %d" % start)
    assert 0.05 < valid.float().mean().item() < 0.6


# ------------------------------------------------------------------ self-filter
def test_filter_points_bit_exact_vs_reference_golden(golden):
    g = golden("self_filter.npz")
    for i in range(int(g["n_cases"])):
        c32 = oin.filter_centres_f32(g[f"q{i}"][None], g[f"base{i}"][None])[0]
        keep = engine.filter_points_mask(_cuda(g[f"pts{i}"]), _cuda(c32), float(g[f"radius{i}"])).cpu().numpy().astype(bool)
        assert np.array_equal(keep, g[f"keep{i}"])                     # bit-exact, band included


def test_filter_robot_points_dropin():
    """Mirrors test/test_robot_filter.py of the reference."""
    from tygrit_b200.fetch_model import fetch_spheres_dict
    from tygrit_b200.robot_filter import filter_robot_points

    T = ofk.link_frames_4x4(np.zeros((1, 10)))[0]
    from oracle.fetch_model import LINK_NAMES

    poses = {n: T[i] for i, n in enumerate(LINK_NAMES)}
    sph = fetch_spheres_dict()
    assert filter_robot_points(np.empty((0, 3), np.float32), poses, sph).shape == (0, 3)
    far = np.array([[5, 5, 1], [10, 0, 0.5], [-3, -3, 2]], np.float32)
    assert filter_robot_points(far, poses, sph).shape[0] == 3
    pts = np.vstack([[[0, 0, 0.2]], [[5, 5, 1]]]).astype(np.float32)
    out = filter_robot_points(pts, poses, sph, sphere_radius=0.15)
    assert out.shape[0] == 1 and out.dtype == np.float32 and np.allclose(out[0], [5, 5, 1])
    arm_pts = np.vstack([poses["gripper_link"][:3, 3], poses["elbow_flex_link"][:3, 3], [5, 5, 1]]).astype(np.float32)
    assert filter_robot_points(arm_pts, poses, sph, sphere_radius=0.12).shape[0] == 1
    big = np.random.default_rng(42).uniform(-5, 5, (10_000, 3)).astype(np.float32)
    n = filter_robot_points(big, poses, sph).shape[0]
    assert 0 < n < 10_000
    t = filter_robot_points(_cuda(big), poses, sph)
    assert t.is_cuda and t.shape[0] == n


def _numpy_depth(states, boxes, H, W, K, seed):
    """numpy twin of csrc/synth.cu (without the invalid-pixel hash): analytic ray casting."""
    from oracle import fetch_model as om

    out = np.zeros((len(states), H, W), np.float32)
    for s, st in enumerate(states.astype(np.float64)):
        cam = ofk.camera_pose(st[None, 3:11], st[None, 11:13], st[None, :3])[0]
        cen = ofk.sphere_centres(st[None, 3:11], st[None, :3], st[None, 11:13])[0]
        v, u = np.mgrid[0:H, 0:W]
        d = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u, float)], -1) @ cam[:3, :3].T
        o = cam[:3, 3]
        with np.errstate(divide="ignore", invalid="ignore"):
            t = np.where(d > 0, (np.array([4.0, 3.0, 2.6]) - o) / d, np.where(d < 0, (np.array([-4.0, -3.0, 0.0]) - o) / d, 1e30))
            best = t.min(-1)
            for lo, hi in boxes:
                ta, tb = (np.array(lo) - o) / d, (np.array(hi) - o) / d
                t0 = np.maximum(np.minimum(ta, tb).max(-1), 0.0)
                t1 = np.minimum(np.maximum(ta, tb).min(-1), best)
                best = np.where((t0 <= t1) & (t0 > 0), t0, best)
            dd = (d * d).sum(-1)
            for c, r in zip(cen, om.SPHERE_LOCAL[:, 3]):
                l = c - o
                bq = d @ l
                disc = bq * bq - dd * (l @ l - r * r)
                tt = (bq - np.sqrt(np.maximum(disc, 0))) / dd
                best = np.where((disc >= 0) & (tt > 0.02) & (tt < best), tt, best)
        out[s] = best
    return out


def test_self_filter_depth_batch_vs_oracle():
    """BASELINE config[2] at oracle-sized scale: batched depth self-filter, FK on the GPU."""
    H, W = 120, 160
    K = synth.HEAD_CAMERA_K.copy()
    K[:2] *= W / 640.0
    states = synth.random_states(6, seed=9)
    states[:, 11] = np.clip(states[:, 11], -0.6, 0.6)
    states[:, 12] = 0.9                                              # look down at the arm
    boxes = synth.room_boxes()
    depth = _numpy_depth(states, boxes, H, W, K, 0)
    depth[:, ::7, ::5] = 0.0
    depth[0, 3, 4] = np.nan
    for stride in (1, 2, 4):
        mask, world = engine.self_filter_depth(_cuda(depth), K, _cuda(states), 0.08, (0.2, 3.0), stride, want_world=True)
        mask, world = mask.cpu().numpy().astype(bool), world.cpu().numpy()
        n_removed = 0
        for s in range(len(states)):
            T = ofk.camera_pose(states[s:s + 1, 3:11], states[s:s + 1, 11:13], states[s:s + 1, :3])[0]
            pw, keep = oin.depth_to_world(depth[s], K, T, (0.2, 3.0), stride)
            np.testing.assert_allclose(world[s][keep], pw, atol=5e-6)   # fp32 camera pose vs fp64 pose
            c64 = ofk.sphere_centres(states[s:s + 1, 3:11], states[s:s + 1, :3], states[s:s + 1, 11:13])[0]
            signed = oin.self_filter_signed_band(pw, c64, 0.08)
            expect = np.zeros(keep.shape, bool)
            expect[keep] = signed >= 0
            decisive = np.ones(keep.shape, bool)
            decisive[keep] = np.abs(signed) > BAND
            assert np.array_equal(mask[s][decisive], expect[decisive])
            assert not mask[s][~keep].any()                             # invalid depth is never kept
            n_removed += int((signed < 0).sum())
        assert n_removed > 50                                           # the robot was really in view
        # the mask-only paths (streaming kernel at stride 1, camera-frame prefilter otherwise) must give the mask of the
        # path that also writes the world points, byte for byte
        fast = engine.self_filter_depth(_cuda(depth), K, _cuda(states), 0.08, (0.2, 3.0), stride).cpu().numpy().astype(bool)
        assert np.array_equal(fast, mask)


@pytest.mark.parametrize("shape", [(480, 640, 24), (100, 164, 24), (480, 640, 96), (37, 2048, 5)])
def test_self_filter_fast_path_equals_generic_kernel(shape):
    """Full-resolution frames (a size with ragged tiles and a short last chunk, enough states for the shared-memory
    ring to wrap several times, and a wide image with 3 rows per chunk): k_self_filter_stream vs k_self_filter_depth."""
    H, W, n_states = shape
    K = synth.HEAD_CAMERA_K.copy()
    K[:2] *= W / 640.0
    states = synth.random_states(n_states, seed=21)
    states[::2, 12] = 0.9                                             # half of the states look down at the arm
    st = _cuda(states)
    bx = _cuda(np.array([[*lo, *hi] for lo, hi in synth.room_boxes()], np.float32))
    depth = torch.empty(len(states), H, W, device="cuda")
    from tygrit_b200 import _native as N
    N.check(N.lib().tg_synth_depth(st.data_ptr(), len(states), bx.data_ptr(), bx.shape[0], H, W, np.ascontiguousarray(K).ctypes.data,
                                   0.0, 3, depth.data_ptr(), None), "tg_synth_depth")
    depth[1, 5, 7] = float("nan")
    depth[2, 9, 11] = float("inf")
    slow, _ = engine.self_filter_depth(depth, K, st, 0.08, (0.2, 3.0), 1, want_world=True)
    fast = engine.self_filter_depth(depth, K, st, 0.08, (0.2, 3.0), 1)
    assert torch.equal(fast, slow)
    for stride in (2, 4):                                             # strided mask-only path vs the path with world output
        slow_s, _ = engine.self_filter_depth(depth, K, st, 0.08, (0.2, 3.0), stride, want_world=True)
        assert torch.equal(engine.self_filter_depth(depth, K, st, 0.08, (0.2, 3.0), stride), slow_s)
    removed = ((depth > 0.2) & (depth < 3.0) & (slow == 0)).float().mean().item()
    assert removed > 1e-3                                             # robot pixels exist in these frames


def test_synth_depth_kernel_matches_numpy_twin():
    from tygrit_b200 import _native as N

    H, W = 60, 80
    K = synth.HEAD_CAMERA_K.copy()
    K[:2] *= W / 640.0
    states = synth.random_states(3, seed=4)
    boxes = synth.room_boxes()
    bx = _cuda(np.array([[*lo, *hi] for lo, hi in boxes], np.float32))
    out = torch.empty(3, H, W, device="cuda")
    st = _cuda(states)
    Kc = np.ascontiguousarray(K)
    N.check(N.lib().tg_synth_depth(st.data_ptr(), 3, bx.data_ptr(), len(boxes), H, W, Kc.ctypes.data, 0.0, 1,
                                   out.data_ptr(), None), "tg_synth_depth")
    got = out.cpu().numpy()
    ref = _numpy_depth(states, boxes, H, W, K, 1)
    ok = got > 0
    assert 0.9 < ok.mean() < 0.995                                    # ~2 % invalid pixels
    assert np.median(np.abs(got[ok] - ref[ok])) < 1e-4 and (np.abs(got[ok] - ref[ok]) < 1e-2).mean() > 0.98


# ------------------------------------------------------------------ the benchmarked configuration itself
@pytest.fixture(scope="module")
def bench_world():
    """What bench.py times: the 500k-point room cloud with the distance grid at 100 voxels/m (bench.py's --vpm default)."""
    cloud = synth.room_cloud(500_000)
    w = CollisionWorld(max_points=1 << 20, voxels_per_metre=100.0, max_voxels=192 << 20)
    w.build(_cuda(cloud))
    return cloud, w, cchecker.CCloud(cloud)


def test_bench_configuration_500k_100vpm_vs_fp64_oracle(bench_world):
    cloud, w, cc = bench_world
    dev = torch.device("cuda", 0)
    arm, base = synth.device_configs(300_000, seed=42, device=dev)       # bench.py's own generator and seed
    valid = w.check_configs(arm, base).cpu().numpy()
    cl = cc.clearance(arm.cpu().numpy(), base.cpu().numpy(), cap=1e-3)
    _assert_verdicts(valid, cl, "bench C4 sample")
    assert 0.1 < valid.mean() < 0.3
    # the planner-like batch of bench.py's free_workload leg
    f_arm, f_base, floor = synth.free_base_configs(w, 200_000, seed=42, device=dev)
    f_valid = w.check_configs(f_arm, f_base).cpu().numpy()
    _assert_verdicts(f_valid, cc.clearance(f_arm.cpu().numpy(), f_base.cpu().numpy(), cap=1e-3), "bench free workload")
    assert f_valid.mean() > 0.5 and 0.05 < floor < 0.9
    # and edges on the same world
    a0, b0, a1, b1 = synth.random_edges(3000, seed=5)
    ev = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1)).cpu().numpy()
    ecl, _ = cc.edge_clearance(a0, b0, a1, b1, cap=1e-3)
    _assert_verdicts(ev, ecl, "bench-world edges")


def test_coarse_far_field_variant_gives_identical_verdicts(bench_world, tabletop):
    """Batches of >= 2^18 configurations run the variant of the configuration kernel that holds the coarse far-field grid
    in shared memory (most spheres of a free configuration are proven free without touching the fine grid).  Its verdicts
    and its ambiguous-item counts must be those of the plain variant, on a large world and on a small one."""
    dev = torch.device("cuda", 0)
    for (cloud, w, cc), box in ((bench_world, (-3.5, 3.5, -2.5, 2.5)), (tabletop, (-1.5, 2.5, -2.0, 2.0))):
        arm, base = synth.device_configs(400_000, seed=77, device=dev, base_box=box)
        f_arm, f_base, _ = synth.free_base_configs(w, 300_000, seed=78, device=dev, base_box=box)
        for a, b in ((arm, base), (f_arm, f_base)):
            w.counters()
            with_coarse = w.check_configs(a, b)
            c1 = w.counters()
            try:
                w.set_coarse(False)
                plain = w.check_configs(a, b)
                c0 = w.counters()
            finally:
                w.set_coarse(True)
            assert torch.equal(with_coarse, plain)
            assert c1["narrow_items"] <= c0["narrow_items"] and c1["definite_collisions"] == c0["definite_collisions"]
            _assert_verdicts(with_coarse[:100_000].cpu().numpy(), cc.clearance(a[:100_000].cpu().numpy(), b[:100_000].cpu().numpy(), cap=1e-3),
                             "coarse variant")


def test_self_collision_term_vs_oracle(tabletop):
    """Optional sphere-vs-sphere term over the reference's SELF_COLLISION_PAIRS (rl/train.py:78-107): off by default,
    on = verdict of min(cloud clearance, self clearance); clearance output follows; edges sample it too."""
    cloud, w, cc = tabletop
    arm, base = synth.random_configs(40_000, seed=23, base_box=(-1, 1, -1, 1))
    cl_cloud = cc.clearance(arm, base)
    cl_self = cchecker.self_clearance(arm)
    np.testing.assert_allclose(cl_self[:2000], collide.self_clearance(arm[:2000]), atol=1e-12)
    assert 0.05 < (cl_self <= 0).mean() < 0.3
    off = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    try:
        w.set_self_collision(True)
        on, clr = w.check_configs(_cuda(arm), _cuda(base), clearance=True)
        on, clr = on.cpu().numpy(), clr.cpu().numpy()
        a0, b0, a1, b1 = synth.random_edges(1500, seed=24, base_box=(-1, 1, -1, 1), rrt_like=True)
        ev = w.check_edges(_cuda(a0), _cuda(b0), _cuda(a1), _cuda(b1)).cpu().numpy()
        # empty world: only the self term is left
        e = CollisionWorld(max_points=1024)
        e.build(np.empty((0, 3), np.float32))
        e.set_self_collision(True)
        only_self = e.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
        tuck = np.array([synth.TUCK_ARM], np.float32)
        assert e.check_configs(_cuda(tuck), _cuda(np.zeros((1, 3), np.float32))).item() == 1   # test/test_planner.py:72-78
    finally:
        w.set_self_collision(False)
    _assert_verdicts(off, cl_cloud, "self off")
    total = np.minimum(cl_cloud, cl_self)
    _assert_verdicts(on, total, "self on")
    _assert_verdicts(only_self, cl_self, "self only")
    assert (off != on).sum() > 100                                      # the term decided something
    np.testing.assert_allclose(clr, np.minimum(total, 0.25), atol=3e-6)
    # edge oracle: min over the samples of min(cloud, self)
    ea, eb = collide.edge_vectors(a0, b0), collide.edge_vectors(a1, b1)
    n = collide.edge_n(ea, eb)
    ecl = np.empty(len(ea))
    for i in range(len(ea)):
        q = collide.edge_samples(ea[i], eb[i], int(n[i])).astype(np.float32)
        ecl[i] = min(cc.clearance(q[:, 3:], q[:, :3]).min(), cchecker.self_clearance(q[:, 3:]).min())
    _assert_verdicts(ev, ecl, "self on, edges")
    assert np.array_equal(w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy(), off)   # and off again


def test_streaming_self_filter_full_resolution_vs_oracle():
    """BASELINE config[2] at its real frame size: k_self_filter_stream (the kernel C3 times) at 480 x 640 compared with
    the oracle directly (oracle/ingest.py = kinematics/robot_filter.py + utils/depth.py), not with the generic kernel."""
    from tygrit_b200 import _native as N

    H, W, S = 480, 640, 3
    K = synth.HEAD_CAMERA_K.copy()
    states = synth.random_states(S, seed=31)
    states[:, 11] = np.clip(states[:, 11], -0.6, 0.6)
    states[:, 12] = 0.9
    st = _cuda(states)
    bx = _cuda(np.array([[*lo, *hi] for lo, hi in synth.room_boxes()], np.float32))
    depth = torch.empty(S, H, W, device="cuda")
    N.check(N.lib().tg_synth_depth(st.data_ptr(), S, bx.data_ptr(), bx.shape[0], H, W, np.ascontiguousarray(K).ctypes.data,
                                   0.0, 5, depth.data_ptr(), None), "tg_synth_depth")
    mask = engine.self_filter_depth(depth, K, st, 0.08, (0.2, 3.0), 1).cpu().numpy().astype(bool)   # mask only, stride 1: streaming kernel
    d_np = depth.cpu().numpy()
    removed = 0
    for s in range(S):
        T = ofk.camera_pose(states[s:s + 1, 3:11], states[s:s + 1, 11:13], states[s:s + 1, :3])[0]
        pw, keep = oin.depth_to_world(d_np[s], K, T, (0.2, 3.0), 1)
        c64 = ofk.sphere_centres(states[s:s + 1, 3:11], states[s:s + 1, :3], states[s:s + 1, 11:13])[0]
        signed = oin.self_filter_signed_band(pw, c64, 0.08)
        expect = np.zeros(keep.shape, bool)
        expect[keep] = signed >= 0
        decisive = np.ones(keep.shape, bool)
        decisive[keep] = np.abs(signed) > BAND
        assert np.array_equal(mask[s][decisive], expect[decisive])
        assert not mask[s][~keep].any()
        removed += int((signed < 0).sum())
    assert removed > 1000


def test_replanning_step_graph_runner():
    """BASELINE config 5 as CUDA-graph replays timed from C (tg_c5_run): fixed bounds required, the world left behind is
    the last cloud's (verdicts vs the oracle), every step validates fresh edges, and a build after it still works."""
    clouds_np = [synth.room_cloud(60_000 + 20_000 * i, perturb_step=i) for i in range(3)]
    w = CollisionWorld(max_points=1 << 18)
    with pytest.raises(RuntimeError):
        w.run_replanning_steps([_cuda(c) for c in clouds_np], 1000, steps=2)       # no fixed bounds
    w.set_bounds((-4.2, -3.2, 0.2), (4.2, 3.2, 2.8))
    r = w.run_replanning_steps([_cuda(c) for c in clouds_np], 20_000, steps=9, warmup=2, seed=3)
    assert r["graphs"] == 6 and r["nodes_per_step_graph"] >= 10
    assert 0.0 < r["build_ms"] < r["ms_per_step"] < 50.0
    assert 0.02 < r["valid_fraction"] < 0.5 and r["config_checks_per_step"] > 20_000 * 8 * 10
    # 9 steps over 3 clouds: the last replay built clouds[2]
    arm, base = synth.random_configs(20_000, seed=12)
    got = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    _assert_verdicts(got, cchecker.CCloud(clouds_np[2]).clearance(arm, base, cap=1e-3), "after graph replays")
    r2 = w.run_replanning_steps([_cuda(c) for c in clouds_np], 20_000, steps=9, warmup=2, seed=4)
    assert abs(r2["valid_fraction"] - r["valid_fraction"]) < 0.05 and r2["valid_fraction"] != r["valid_fraction"]   # other edges
    w.build(_cuda(clouds_np[0]))
    got = w.check_configs(_cuda(arm), _cuda(base)).cpu().numpy()
    _assert_verdicts(got, cchecker.CCloud(clouds_np[0]).clearance(arm, base, cap=1e-3), "rebuild after graph replays")


def test_verdict_bit_packing_round_trip():
    """tg_pack_bits / tg_unpack_bits (the gather of a sharded job moves verdict bits): ragged sizes, unaligned views."""
    g = torch.Generator(device="cuda").manual_seed(3)
    for n in (1, 7, 8, 9, 1000, 65_537, 1 << 20):
        v = (torch.rand(n + 3, generator=g, device="cuda") < 0.3).to(torch.uint8) * 255   # any non-zero byte is "valid"
        for view in (v[:n], v[3:3 + n]):                                                   # second view: not 8-byte aligned
            bits = engine.pack_verdicts(view)
            assert bits.shape[0] == (n + 7) // 8
            ref = np.packbits((view.cpu().numpy() != 0).astype(np.uint8), bitorder="little")
            assert np.array_equal(bits.cpu().numpy(), ref)
            assert torch.equal(engine.unpack_verdicts(bits, n), (view != 0).to(torch.uint8))
