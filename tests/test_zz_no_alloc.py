"""GPU: ``tg_check_configs_noalloc`` — the configuration check with its queue in a static device array of the library, for
callers that must not allocate (include/tygrit_b200.h; no counterpart in the reference).
Called through the C ABI and compared with ``tg_check_configs`` on the same inputs: the verdict bytes must be
identical (same kernels, same arithmetic; the bar of every verdict test here is bit-exact), on two calls in a row — the
second sees the first call's leftovers in the shared array — and on both kernel variants (below / from 2^18 configurations
per call).  Collected last (file name) so that a failure here hides no other test; passes alone in a fresh process too.
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _world():
    from tygrit_b200 import synth
    from tygrit_b200.engine import CollisionWorld

    cloud = synth.room_cloud(120_000)
    w = CollisionWorld(device=0, max_points=1 << 18)
    w.build(torch.from_numpy(cloud).cuda())
    return w, cloud


def _noalloc(w, arm, base, valid):
    from tygrit_b200 import _native as N

    stream = torch.cuda.current_stream(w.device).cuda_stream
    with torch.cuda.device(w.device):
        return N.lib().tg_check_configs_noalloc(w._h, arm.data_ptr(), base.data_ptr(), arm.shape[0], valid.data_ptr(), stream)


def test_noalloc_matches_check_configs_twice_in_a_row(built_lib):
    from oracle import cchecker
    from tygrit_b200 import _native as N, synth

    w, cloud = _world()
    dev = torch.device("cuda", 0)
    batches = []
    for seed, n in ((11, 100_003), (12, N.TG_NOALLOC_MAX_CONFIGS)):      # plain kernel, then the coarse-grid variant at the limit
        arm, base = synth.random_configs(n, seed=seed)
        batches.append((torch.from_numpy(arm).to(dev), torch.from_numpy(base).to(dev), arm, base))
    expected = [w.check_configs(a, b).clone() for a, b, _, _ in batches]
    torch.cuda.synchronize()
    outs = [torch.full((a.shape[0],), 7, dtype=torch.uint8, device=dev) for a, _, _, _ in batches]   # allocated BEFORE the calls
    free0 = torch.cuda.mem_get_info(0)[0]
    for rep in range(2):                                                  # two calls in a row per batch, interleaved
        for (a, b, _, _), out in zip(batches, outs):
            out.fill_(7)
            N.check(_noalloc(w, a, b, out), "tg_check_configs_noalloc")
        torch.cuda.synchronize()
        for k, (out, exp) in enumerate(zip(outs, expected)):
            assert torch.equal(out, exp), f"repetition {rep}, batch {k}: {int((out != exp).sum())} verdicts differ from tg_check_configs"
    assert torch.cuda.mem_get_info(0)[0] >= free0, "device memory was allocated during the no-alloc calls"
    # and both agree with the fp64 oracle outside the band
    _, _, arm, base = batches[0]
    cl = cchecker.CCloud(cloud).clearance(arm[:50_000], base[:50_000], cap=1e-3)
    d = np.abs(cl) > 1e-5
    assert np.array_equal(outs[0][:50_000].cpu().numpy().astype(bool)[d], (cl > 0)[d])


def test_noalloc_small_queue_and_self_collision(built_lib):
    """The inline-overflow path (queue limit hook) and the optional self-collision term give tg_check_configs' bytes too."""
    from tygrit_b200 import _native as N, synth

    w, _ = _world()
    dev = torch.device("cuda", 0)
    arm, base = synth.random_configs(40_000, seed=13)
    a, b = torch.from_numpy(arm).to(dev), torch.from_numpy(base).to(dev)
    out = torch.empty(a.shape[0], dtype=torch.uint8, device=dev)
    w.set_self_collision(True)
    N.check(N.lib().tg_env_set_queue_limit(w._h, 64), "tg_env_set_queue_limit")
    try:
        exp = w.check_configs(a, b)
        for _ in range(2):
            N.check(_noalloc(w, a, b, out), "tg_check_configs_noalloc")
            torch.cuda.synchronize()
            assert torch.equal(out, exp)
    finally:
        N.check(N.lib().tg_env_set_queue_limit(w._h, 0), "tg_env_set_queue_limit")
        w.set_self_collision(False)


def test_noalloc_rejects_oversize_and_accepts_empty(built_lib):
    from tygrit_b200 import _native as N

    w, _ = _world()
    dev = torch.device("cuda", 0)
    n = N.TG_NOALLOC_MAX_CONFIGS + 1
    arm = torch.zeros((n, 8), dtype=torch.float32, device=dev)
    base = torch.zeros((n, 3), dtype=torch.float32, device=dev)
    out = torch.full((n,), 7, dtype=torch.uint8, device=dev)
    rc = _noalloc(w, arm, base, out)
    assert rc == N.TG_E_CAPACITY
    msg = N.lib().tg_last_error().decode()
    assert "exceed the maximum" in msg and str(N.TG_NOALLOC_MAX_CONFIGS) in msg
    torch.cuda.synchronize()
    assert bool((out == 7).all()), "a rejected call must not touch the output"
    assert _noalloc(w, arm[:0], base[:0], out[:0]) == 0
    with pytest.raises(RuntimeError):
        N.check(rc, "tg_check_configs_noalloc")
