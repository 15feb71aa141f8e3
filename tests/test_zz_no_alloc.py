"""rgbdfe_match_knn2_noalloc (no counterpart in the reference; static __device__ scratch, no cudaMalloc) against
rgbdfe_match_knn2 through the C ABI: bit-exact match lists (the tolerance of every match test here), two calls in a
row with different sizes so that the second sees the first call's leftovers in the shared scratch.  Collected last
on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(rng, nq, nt):
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    k = min(nq, nt) // 2
    t[:k] = q[:k]  # true correspondences with a few flipped bits
    t[:k, 0] ^= rng.integers(0, 8, k, dtype=np.uint8)
    if nt > 70:
        t[40:60] = t[40]  # exact ties: lower train index must win
    q_xyz = rng.uniform(0.5, 3.0, (nq, 3)).astype(np.float32)
    t_xyz = rng.uniform(0.5, 3.0, (nt, 3)).astype(np.float32)
    t_xyz[::9, 2] = 0  # invalid depth
    outl = (rng.random(nq) < 0.15).astype(np.uint8)
    return q, t, outl, q_xyz, t_xyz


def _noalloc(ctx, torch, case, ratio, discard, cap=None):
    import rgbdfe

    q, t, outl, q_xyz, t_xyz = case
    dev = [torch.from_numpy(a).cuda() for a in (q, t, outl, q_xyz, t_xyz)]
    cap = len(q) if cap is None else cap
    out = torch.zeros(max(cap, 1) * 16, dtype=torch.uint8, device="cuda")
    n = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    vp = lambda x: C.c_void_p(x.data_ptr())
    rc = rgbdfe.lib().rgbdfe_match_knn2_noalloc(ctx.h, None, vp(dev[0]), len(q), vp(dev[1]), len(t), vp(dev[2]), vp(dev[3]), vp(dev[4]),
                                                C.c_float(ratio), int(discard), vp(out), cap, vp(n))
    ctx._chk(rc)
    ctx.sync()
    cnt = int(n.cpu()[0])
    return out.cpu().numpy().view(rgbdfe.DM_DTYPE)[:cnt].copy()


def test_noalloc_matches_existing_entry_twice_in_a_row():
    import torch

    import rgbdfe

    ctx = rgbdfe.Context(0, 640, 480, 4096 - 64, 1)
    try:
        rng = np.random.default_rng(77)
        # big first, small second: the second call must not see the first call's partials or counts
        for nq, nt, ratio, discard in ((2000, 1873, 0.9, True), (301, 517, 0.75, False), (4096, 4096, 0.9, True), (130, 1, 0.9, True)):
            case = _case(rng, nq, nt)
            want = ctx.match(case[0], case[1], case[2], case[3], case[4], ratio, discard)
            got = _noalloc(ctx, torch, case, ratio, discard)
            assert len(want) == len(got) and (want == got).all(), (nq, nt)
            if nt > 1:
                assert len(want) > 0
        # cap smaller than the result: the first `cap` matches, *n = cap
        case = _case(rng, 600, 600)
        want = ctx.match(case[0], case[1], None, case[3], case[4], 0.9, False)
        got = _noalloc(ctx, torch, case, 0.9, False, cap=50)
        assert len(want) > 50 and len(got) == 50 and (want[:50] == got).all()
        # empty sides
        assert len(_noalloc(ctx, torch, _case(rng, 0, 50), 0.9, True, cap=4)) == 0
    finally:
        ctx.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    import rgbdfe

    ctx = rgbdfe.Context(0, 640, 480, 1000, 1)
    try:
        with pytest.raises(rgbdfe.RgbdfeError, match="4096"):
            _noalloc(ctx, torch, _case(np.random.default_rng(1), 4097, 10), 0.9, True)
    finally:
        ctx.close()
