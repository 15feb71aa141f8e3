"""Parity of the CUDA path (through the C ABI) against the CPU oracle, stage by stage.

Bar (north_star): keypoints, descriptors, match indices and RANSAC inlier sets bit-exact;
poses within 1e-5 (rotation entries and translation, FP32).
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-5


@pytest.fixture(scope="module")
def ctx():
    import rgbdfe

    c = rgbdfe.Context(0, 640, 480, 2000, 4)
    yield c
    c.close()


def _unproject(cam, kps, depth):
    import synth

    return synth.unproject(cam, np.stack([kps["x"], kps["y"]], 1), depth)


@pytest.mark.parametrize("nfeat", [1000, 2000, 300])
def test_orb_stages_bit_exact(ctx, oracle, seq_vga, nfeat):
    cam, frames = seq_vga
    ctx.orb_configure(nfeat, 1.2, 8, 20, 7)
    orb = oracle.Orb(nfeat)
    t_o, t_g = orb.tables(), ctx.orb_tables()
    for k in ("scale", "inv_scale", "sigma2", "inv_sigma2", "quota"):
        assert (t_o[k] == t_g[k]).all(), k
    for gray, _ in frames[:2]:
        kps_o, desc_o = orb.extract(gray)
        kps_g, desc_g = ctx.orb_extract(gray)
        for l in range(8):
            assert (ctx.pyramid_level(l) == orb.level_padded(l)).all(), f"pyramid level {l}"
            c_o, c_g = orb.level_candidates(l), ctx.stage_candidates(l)
            assert len(c_o) == len(c_g), f"FAST candidate count level {l}: {len(c_o)} vs {len(c_g)}"
            for f in ("x", "y", "response"):
                assert (c_o[f] == c_g[f]).all(), f"FAST candidates level {l} field {f}"
            b = orb.level_blurred(l)
            if b is not None:
                assert (ctx.stage_blurred(l) == b).all(), f"blur level {l}"
        assert len(kps_o) == len(kps_g)
        for f in ("x", "y", "size", "response", "octave", "class_id"):
            assert (kps_o[f] == kps_g[f]).all(), f"keypoint field {f}"
        assert (kps_o["angle"].view(np.uint32) == kps_g["angle"].view(np.uint32)).all(), "angle bit pattern"
        assert (desc_o == desc_g).all(), "descriptors"


@pytest.mark.parametrize("size", [(611, 397), (322, 251), (1000, 80)])
def test_orb_odd_image_sizes(oracle, seq_vga, size):
    """Widths and heights that are not multiples of 4 or 16 (row pitches with padding, level-0 spans that straddle the ring,
    row groups that end inside a group of four): pyramid bytes with ring, keypoints and descriptors against the oracle."""
    import rgbdfe

    w, h = size
    _, frames = seq_vga
    big = np.concatenate([np.concatenate([frames[0][0], frames[1][0]], 1), np.concatenate([frames[2][0], frames[3][0]], 1)], 0)
    gray = np.ascontiguousarray(big[37 : 37 + h, 91 : 91 + w])
    ctx = rgbdfe.Context(0, w, h, 500, 1)
    try:
        nlev = 8 if h >= 200 else 2  # every level must stay above 40 px
        ctx.orb_configure(500, 1.2, nlev, 20, 7)
        orb = oracle.Orb(500, 1.2, nlev)
        kps_o, desc_o = orb.extract(gray)
        kps_g, desc_g = ctx.orb_extract(gray)
        for l in range(nlev):
            assert (ctx.pyramid_level(l) == orb.level_padded(l)).all(), f"pyramid level {l}"
        assert len(kps_o) == len(kps_g) and len(kps_o) > 50
        for f in ("x", "y", "response", "octave"):
            assert (kps_o[f] == kps_g[f]).all(), f
        assert (kps_o["angle"].view(np.uint32) == kps_g["angle"].view(np.uint32)).all()
        assert (desc_o == desc_g).all()
    finally:
        ctx.close()


def test_orb_blur_on_side_stream_same_result(ctx, oracle, seq_vga):
    """Option overlap_blur: K3 beside K4 on a side stream; several back-to-back extractions, same keypoints and descriptors."""
    cam, frames = seq_vga
    ctx.orb_configure(1000, 1.2, 8, 20, 7)
    orb = oracle.Orb(1000)
    ctx.set_option("overlap_blur", 1)
    try:
        for gray, _ in frames[:3]:
            kps_o, desc_o = orb.extract(gray)
            kps_g, desc_g = ctx.orb_extract(gray)
            assert len(kps_o) == len(kps_g)
            for f in ("x", "y", "response", "octave"):
                assert (kps_o[f] == kps_g[f]).all(), f
            assert (kps_o["angle"].view(np.uint32) == kps_g["angle"].view(np.uint32)).all()
            assert (desc_o == desc_g).all()
    finally:
        ctx.set_option("overlap_blur", 0)


def test_contexts_sharing_one_stream(ctx, oracle, seq_vga):
    """rgbdfe_share_stream: a second context enqueues on the first one's stream; interleaved extractions stay exact."""
    import rgbdfe

    cam, frames = seq_vga
    other = rgbdfe.Context(0, 640, 480, 1000, 2)
    try:
        other.share_stream(ctx)
        assert other.stream == ctx.stream
        ctx.orb_configure(1000, 1.2, 8, 20, 7)
        other.orb_configure(500, 1.2, 8, 20, 7)
        o1, o2 = oracle.Orb(1000), oracle.Orb(500)
        for gray, _ in frames[:2]:
            k1, d1 = ctx.orb_extract(gray)
            k2, d2 = other.orb_extract(gray)
            ko1, do1 = o1.extract(gray)
            ko2, do2 = o2.extract(gray)
            assert len(k1) == len(ko1) and (d1 == do1).all() and (k1["x"] == ko1["x"]).all()
            assert len(k2) == len(ko2) and (d2 == do2).all() and (k2["x"] == ko2["x"]).all()
    finally:
        other.close()


def test_shared_stream_outlives_its_owner(oracle, seq_vga):
    """The owner of a shared stream is destroyed first: the stream passes to a sharer and keeps working."""
    import rgbdfe

    cam, frames = seq_vga
    owner = rgbdfe.Context(0, 640, 480, 1000, 1)
    a, b = rgbdfe.Context(0, 640, 480, 1000, 1), rgbdfe.Context(0, 640, 480, 1000, 1)
    try:
        a.share_stream(owner)
        b.share_stream(a)  # sharing a sharer's stream = sharing its owner's
        s = owner.stream
        assert a.stream == s and b.stream == s
        owner.close()
        ko, do = oracle.Orb(1000).extract(frames[0][0])
        for c in (a, b):
            k, d = c.orb_extract(frames[0][0])
            assert len(k) == len(ko) and (d == do).all()
        a.close()  # the heir goes too: b inherits
        k, d = b.orb_extract(frames[1][0])
        assert len(k) > 100
    finally:
        for c in (owner, a, b):
            c.close()


def test_contexts_on_two_devices_in_one_process(oracle, seq_vga):
    """Kernel attributes (dynamic shared memory limits) are per device: a second context on another GPU of the same
    process must run the large-smem kernels (quadtree, RANSAC prepare / hypotheses) as well."""
    import torch

    import rgbdfe

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cam, frames = seq_vga
    p1, p2, m = _pair(oracle, cam, frames, 0, 1)
    res = []
    for dev in (0, 1):
        c = rgbdfe.Context(dev, 640, 480, 2000, 1)
        try:
            c.orb_configure(2000, 1.2, 8, 20, 7)
            k, d = c.orb_extract(frames[0][0])
            c.rng_seed(42)
            res.append((k, d, c.ransac_se3(p1, p2, m, 200, 10, 3.0, 4)))
        finally:
            c.close()
    assert (res[0][0] == res[1][0]).all() and (res[0][1] == res[1][1]).all()
    _check_se3(res[0][2], res[1][2])


def test_orb_fast_threshold_fallback(ctx, oracle):
    """Low-contrast image: cells with no FAST corner at 20 but some at 7 (ORBextractor.cpp:658-661)."""
    rng = np.random.default_rng(11)
    img = (128 + rng.integers(-9, 10, (480, 640))).astype(np.uint8)
    img[100:300, 200:400] = rng.integers(0, 256, (200, 200), dtype=np.uint8)
    ctx.orb_configure(1000, 1.2, 8, 20, 7)
    orb = oracle.Orb(1000)
    k_o, d_o = orb.extract(img)
    k_g, d_g = ctx.orb_extract(img)
    assert len(k_o) == len(k_g) and (k_o == k_g).all() and (d_o == d_g).all()


def test_orb_flat_image_has_no_keypoints(ctx, oracle):
    img = np.full((480, 640), 77, np.uint8)
    ctx.orb_configure(1000, 1.2, 8, 20, 7)
    k_g, d_g = ctx.orb_extract(img)
    k_o, _ = oracle.Orb(1000).extract(img)
    assert len(k_g) == 0 and len(k_o) == 0


def test_match_bit_exact(ctx, oracle, seq_vga):
    cam, frames = seq_vga
    ctx.orb_configure(1000, 1.2, 8, 20, 7)
    orb = oracle.Orb(1000)
    data = []
    for g, d in frames[:2]:
        k, de = orb.extract(g)
        data.append((k, de, _unproject(cam, k, d)))
    rng = np.random.default_rng(5)
    outl = (rng.random(len(data[0][0])) < 0.2).astype(np.uint8)
    for discard in (True, False):
        for ratio in (0.9, 0.6):
            m_o = oracle.match(data[0][1], data[1][1], outl, data[0][2], data[1][2], ratio, discard)
            m_g = ctx.match(data[0][1], data[1][1], outl, data[0][2], data[1][2], ratio, discard)
            assert len(m_o) == len(m_g) and (m_o == m_g).all()


def test_match_ties_and_degenerate(ctx, oracle):
    rng = np.random.default_rng(6)
    q = rng.integers(0, 256, (257, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    t[100:140] = t[100]  # exact ties: lower train index must win
    q[:50] = t[100]
    t[200] = q[60]
    xyz_q = np.ones((257, 3), np.float32)
    xyz_t = np.ones((300, 3), np.float32)
    xyz_t[::7, 2] = 0  # invalid depth
    z = np.zeros(257, np.uint8)
    for ratio in (0.9, 1.5):
        m_o = oracle.match(q, t, z, xyz_q, xyz_t, ratio, True)
        m_g = ctx.match(q, t, z, xyz_q, xyz_t, ratio, True)
        assert len(m_o) == len(m_g) and (m_o == m_g).all()
    # < 2 train descriptors: no match (guarded UB of Matcher.cpp:117); empty inputs
    assert len(ctx.match(q, t[:1], z, xyz_q, xyz_t[:1], 0.9, True)) == 0
    assert len(ctx.match(q[:0], t, z[:0], xyz_q[:0], xyz_t, 0.9, True)) == 0


@pytest.mark.parametrize("umma", [2, 1, 0])
def test_match_tile_edges_both_knn_kernels(ctx, oracle, umma):
    """Set sizes around the 128-wide tiles of the tcgen05 kernel (and the 256 / 128 of the mma.sync one), clustered
    descriptors (small distances, many equal ones: the (distance, trainIdx) order decides), several ratios."""
    ctx.set_option("knn_umma", umma)
    try:
        rng = np.random.default_rng(60 + umma)
        for nq, nt in ((1, 2), (127, 129), (128, 128), (129, 127), (255, 257), (300, 1153), (1000, 2), (2000, 1999)):
            centers = rng.integers(0, 256, (24, 32), dtype=np.uint8)
            def cloud(n):
                d = centers[rng.integers(0, len(centers), n)].copy()
                flips = rng.integers(0, 256, (n, 3))  # up to 3 flipped bits per descriptor
                for k in range(3):
                    d[np.arange(n), flips[:, k] >> 3] ^= (1 << (flips[:, k] & 7)).astype(np.uint8)
                return d
            q, t = cloud(nq), cloud(nt)
            xyz_q = np.ones((nq, 3), np.float32)
            xyz_t = np.ones((nt, 3), np.float32)
            z = np.zeros(nq, np.uint8)
            for ratio in (0.7, 0.95, 4.0):
                m_o = oracle.match(q, t, z, xyz_q, xyz_t, ratio, True)
                m_g = ctx.match(q, t, z, xyz_q, xyz_t, ratio, True)
                assert len(m_o) == len(m_g) and (m_o == m_g).all(), (nq, nt, ratio)
    finally:
        ctx.set_option("knn_umma", 1)


def _pair(oracle, cam, frames, i, j, nfeat=1000):
    orb = oracle.Orb(nfeat)
    out = []
    for g, d in (frames[i], frames[j]):
        k, de = orb.extract(g)
        out.append((k, de, _unproject(cam, k, d)))
    m = oracle.match(out[0][1], out[1][1], np.zeros(len(out[0][0]), np.uint8), out[0][2], out[1][2], 0.9, True)
    return out[0][2], out[1][2], m


def _check_se3(r_o, r_g):
    assert r_o["ok"] == r_g["ok"]
    assert r_o["iters_executed"] == r_g["iters_executed"]
    assert len(r_o["inliers"]) == len(r_g["inliers"])
    assert (r_o["inliers"] == r_g["inliers"]).all(), "inlier set / order"
    assert np.abs(r_o["T"] - r_g["T"]).max() <= POSE_TOL
    assert (r_o["T"] == r_g["T"]).all(), "pose bits"
    assert r_o["rmse"] == r_g["rmse"]


def test_ransac_se3_parity_and_rng_stream(ctx, oracle, seq_vga):
    cam, frames = seq_vga
    p1, p2, m = _pair(oracle, cam, frames, 0, 1)
    rng = oracle.Rng(42)
    latch = oracle.Latch()
    ctx.rng_seed(42)
    ctx.depth_cov_set(0, 0.0)
    for it in range(3):  # consecutive calls share the rand() stream and the latch
        r_o = oracle.ransac_se3(p1, p2, m, rng, latch, 200, 10, 3.0, 4)
        r_g = ctx.ransac_se3(p1, p2, m, 200, 10, 3.0, 4)
        _check_se3(r_o, r_g)
        st = ctx.rng_get_state()
        probe = oracle.Rng(1)
        probe.buf[:] = rng.buf
        ctx_next = _next_from_state(st)
        assert ctx_next == probe.next(), "rand() stream position after compute()"
    l, cov = ctx.depth_cov_get()
    assert l == latch.latched.value == 1 and cov == latch.cov.value


def _next_from_state(st):
    r, idx = st[:34].astype(np.uint32), int(st[34])
    return int((int(r[(idx - 31) % 34]) + int(r[(idx - 3) % 34])) & 0xFFFFFFFF) >> 1


def test_ransac_se3_hard_cases(ctx, oracle, seq_vga):
    """Outlier-heavy matches (many iterations, n += 10 skips), too few matches, identity fallback."""
    cam, frames = seq_vga
    p1, p2, m = _pair(oracle, cam, frames, 0, 3)
    rs = np.random.default_rng(9)
    for frac, seed in ((0.5, 1), (0.7, 2), (0.9, 3)):
        mm = m.copy()
        bad = rs.random(len(mm)) < frac
        mm["trainIdx"][bad] = rs.integers(0, len(p2), bad.sum())
        for iters, min_inl in ((200, 10), (500, 30)):
            rng, latch = oracle.Rng(seed), oracle.Latch()
            r_o = oracle.ransac_se3(p1, p2, mm, rng, latch, iters, min_inl, 3.0, 4)
            # the tuning knobs never change results: speculative first chunk 32 / 4 / 1, 512-thread or register-capped small CTAs
            for small, chunk in ((0, 32), (1, 32), (1, 1), (0, 4)):
                ctx.set_option("ransac_small_ctas", small)
                ctx.set_option("ransac_first_chunk", chunk)
                ctx.rng_seed(seed)
                ctx.depth_cov_set(0, 0.0)
                r_g = ctx.ransac_se3(p1, p2, mm, iters, min_inl, 3.0, 4)
                _check_se3(r_o, r_g)
    ctx.set_option("ransac_small_ctas", 0)
    ctx.set_option("ransac_first_chunk", 32)
    # fewer matches than the inlier threshold -> false, identity, stream untouched
    ctx.rng_seed(5)
    before = ctx.rng_get_state()
    r_g = ctx.ransac_se3(p1, p2, m[:5], 200, 10, 3.0, 4)
    assert not r_g["ok"] and len(r_g["inliers"]) == 0 and (r_g["T"] == np.eye(4, dtype=np.float32)).all()
    assert (before == ctx.rng_get_state()).all()
    # identical frames: every hypothesis is perfect; all-equal-distance ties exercise the sort
    mm = m.copy()
    mm["distance"] = 3.0
    rng, latch = oracle.Rng(7), oracle.Latch()
    ctx.rng_seed(7)
    ctx.depth_cov_set(0, 0.0)
    _check_se3(oracle.ransac_se3(p1, p2, mm, rng, latch), ctx.ransac_se3(p1, p2, mm))


@pytest.mark.timeout(240)
def test_ransac_hard_pairs_do_not_need_co_resident_ctas(oracle, seq_vga):
    """A hard pair (no early break) walks all iterations with `ransac_first_chunk` CTAs that take turns; the CTAs of a grid
    need not be co-resident: 24 contexts on 24 host threads at first_chunk = 32 (768 CTAs of 512 threads on 296 slots), and
    one context on a small latency partition.  Results must still equal the oracle's."""
    import threading

    import rgbdfe

    cam, frames = seq_vga
    p1, p2, m = _pair(oracle, cam, frames, 0, 3)
    rs = np.random.default_rng(19)
    mm = m.copy()
    bad = rs.random(len(mm)) < 0.85
    mm["trainIdx"][bad] = rs.integers(0, len(p2), bad.sum())
    want = []
    for seed in (1, 2):
        rng, latch = oracle.Rng(seed), oracle.Latch()
        want.append(oracle.ransac_se3(p1, p2, mm, rng, latch, 500, 30, 3.0, 4))
    assert min(w["iters_executed"] for w in want) > 100  # really a long loop
    n_ctx = 24
    ctxs = [rgbdfe.Context(0, 640, 480, 1000, 1) for _ in range(n_ctx)]
    errs = []

    def work(c, k):
        try:
            c.set_option("ransac_first_chunk", 32)
            for rep in range(3):
                seed = 1 + (k + rep) % 2
                c.rng_seed(seed)
                c.depth_cov_set(0, 0.0)
                _check_se3(want[seed - 1], c.ransac_se3(p1, p2, mm, 500, 30, 3.0, 4))
        except Exception as ex:  # noqa: BLE001
            errs.append(repr(ex))

    try:
        th = [threading.Thread(target=work, args=(c, k)) for k, c in enumerate(ctxs)]
        [t.start() for t in th]
        [t.join() for t in th]
        assert not errs, errs[:3]
    finally:
        for c in ctxs:
            c.close()
    # the same pair on an 8-SM latency partition (16 CTA slots for 32 CTAs); the SM split is fixed per process, so a child does it
    import pickle
    import subprocess
    import sys
    import tempfile

    with tempfile.TemporaryDirectory() as td:
        pickle.dump((p1, p2, mm), open(os.path.join(td, "in.pkl"), "wb"))
        code = (
            "import sys, pickle; sys.path[:0] = %r\n"
            "import rgbdfe\n"
            "p1, p2, mm = pickle.load(open(%r, 'rb'))\n"
            "part = rgbdfe.device_partition(0, 8)\n"
            "c = rgbdfe.Context(0, 640, 480, 1000, 1)\n"
            "if part is not None: c.bind_partition(rgbdfe.PART_LATENCY)\n"
            "c.set_option('ransac_first_chunk', 32); c.rng_seed(1); c.depth_cov_set(0, 0.0)\n"
            "r = c.ransac_se3(p1, p2, mm, 500, 30, 3.0, 4)\n"
            "pickle.dump((part, r), open(%r, 'wb'))\n"
        ) % ([p for p in sys.path if p], os.path.join(td, "in.pkl"), os.path.join(td, "out.pkl"))
        subprocess.run([sys.executable, "-c", code], check=True, timeout=120)
        part, r = pickle.load(open(os.path.join(td, "out.pkl"), "rb"))
        _check_se3(want[0], r)


def test_ransac_umeyama_parity(ctx, oracle, seq_vga):
    """K12 — class Ransac (Solver/Ransac.cpp): inlier list (order), pose bits, executed iterations, rand() position."""
    cam, frames = seq_vga
    p1, p2, m = _pair(oracle, cam, frames, 0, 2)
    rs = np.random.default_rng(11)
    cases = []
    for frac in (0.0, 0.4, 0.8):
        mm = m.copy()
        bad = rs.random(len(mm)) < frac
        mm["trainIdx"][bad] = rs.integers(0, len(p2), bad.sum())
        cases.append(mm)
    seed = 100
    for mm in cases:
        for ev, th, used in ((0, 0.02, 3), (4, 0.01, 3), (0, 0.05, 4)):
            seed += 1
            rng = oracle.Rng(seed)
            ctx.rng_seed(seed)
            r_o = oracle.ransac_umeyama(p1, p2, mm, rng, ev, th, 0.2, 10, used)
            r_g = ctx.ransac_umeyama(p1, p2, mm, ev, th, 0.2, 10, used)
            assert r_o["ok"] == r_g["ok"]
            assert r_o["iters_executed"] == r_g["iters_executed"]
            assert len(r_o["inliers"]) == len(r_g["inliers"]) and (r_o["inliers"] == r_g["inliers"]).all()
            assert np.abs(r_o["T"] - r_g["T"]).max() <= POSE_TOL
            assert (r_o["T"] == r_g["T"]).all(), "pose bits"
            probe = oracle.Rng(1)
            probe.buf[:] = rng.buf
            assert _next_from_state(ctx.rng_get_state()) == probe.next(), "rand() stream position after compute()"
    # too few matches after the depth gate -> identity, no inliers, stream untouched (Ransac.cpp:86-89)
    ctx.rng_seed(3)
    before = ctx.rng_get_state()
    r_g = ctx.ransac_umeyama(p1, p2, m[:5], 0, 0.02, 0.2, 10, 3)
    assert not r_g["ok"] and len(r_g["inliers"]) == 0 and (r_g["T"] == np.eye(4, dtype=np.float32)).all()
    assert (before == ctx.rng_get_state()).all()
    # inlier ratio below the gate -> identity (Ransac.cpp:168-171)
    mm = m.copy()
    mm["trainIdx"] = rs.integers(0, len(p2), len(mm))
    rng = oracle.Rng(9)
    ctx.rng_seed(9)
    r_o = oracle.ransac_umeyama(p1, p2, mm, rng, 0, 0.02, 0.5, 10, 3)
    r_g = ctx.ransac_umeyama(p1, p2, mm, 0, 0.02, 0.5, 10, 3)
    assert r_o["ok"] == r_g["ok"] and r_o["iters_executed"] == r_g["iters_executed"] and len(r_g["inliers"]) == len(r_o["inliers"])
    assert (r_o["T"] == r_g["T"]).all()
    # the error versions that crash in the reference are refused, not emulated
    import rgbdfe

    with pytest.raises(rgbdfe.RgbdfeError):
        ctx.ransac_umeyama(p1, p2, m, 1, 0.02, 0.2, 10, 3)


@pytest.mark.parametrize("chain", [False, True, "partitioned"])
def test_tracking_pipeline_device_resident(oracle, seq_vga, chain):
    """T2: extract (batched) -> match -> RansacSE3 with outlier-flag carry-over, frame to frame;
    pair by pair (rgbdfe_track), as one rgbdfe_track_sequence call (batched kNN + fused filter), and the same
    call from a context bound to the latency SM partition (kNN on the throughput partition)."""
    import rgbdfe

    cam, frames = seq_vga
    ctx = rgbdfe.Context(0, 640, 480, 1000, 4)
    try:
        if chain == "partitioned":
            part = rgbdfe.device_partition(0, 32)
            if part is None:
                pytest.skip("driver without green contexts")
            assert part[0] >= 32 and part[1] > 0
            ctx.bind_partition(rgbdfe.PART_LATENCY)
        ctx.rng_seed(42)
        fr = [rgbdfe.Frame(ctx) for _ in frames]
        ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
        if chain:
            ctx.track_sequence(fr, 0)
        else:
            for i in range(1, len(frames)):
                ctx.track(fr[i - 1], fr[i], i - 1)
        res = ctx.track_results(0, len(frames) - 1)
        orb = oracle.Orb(1000)
        rng, latch = oracle.Rng(42), oracle.Latch()
        prev = None
        for i, (g, d) in enumerate(frames):
            k, de = orb.extract(g)
            p3 = _unproject(cam, k, d)
            outl = np.zeros(len(k), np.uint8)
            if prev is not None:
                m = oracle.match(prev[1], de, prev[3], prev[2], p3, 0.9, True)
                r = oracle.ransac_se3(prev[2], p3, m, rng, latch, 200, 10, 3.0, 4)
                if len(m) >= 10:
                    outl[m["trainIdx"]] = 1
                if r["ok"]:
                    outl[r["inliers"]["trainIdx"]] = 0
                g_r = res[i - 1]
                assert g_r["n_matches"] == len(m)
                assert bool(g_r["ok"]) == r["ok"] and g_r["n_inliers"] == len(r["inliers"])
                if chain is True:  # mvInliers of every pair of the chain (rgbdfe_track_results_inliers), order included
                    if i == 1:
                        res2, inl_lists = ctx.track_results_inliers(0, len(frames) - 1)
                        assert res2.tobytes() == res.tobytes()
                    assert inl_lists[i - 1].tobytes() == r["inliers"].tobytes()
                assert g_r["iters_executed"] == r["iters_executed"]
                assert np.abs(g_r["T21"].reshape(4, 4) - r["T"]).max() <= POSE_TOL
            gk, gd, gp, go = fr[i].download()
            assert (gk == k).all() and (gd == de).all() and (gp == p3).all()
            assert (go == outl).all(), f"outlier flags of frame {i}"
            prev = (k, de, p3, outl)
        # the batch download (what Frame keeps on the host: mvKeys, mDescriptors, mvKeys3Dc) equals the per-frame one
        cap, nf = ctx.kp_cap, len(fr)
        kps, desc = np.zeros((nf, cap), rgbdfe.KP_DTYPE), np.zeros((nf, cap, 32), np.uint8)
        xyz, n = np.zeros((nf, cap, 3), np.float32), np.zeros(nf, np.int32)
        ctx.frames_download_async(fr, kps, desc, xyz, n, cap)
        ctx.sync()
        for i, f in enumerate(fr):
            gk, gd, gp, _ = f.download()
            assert n[i] == len(gk) and (kps[i, : n[i]] == gk).all() and (desc[i, : n[i]] == gd).all() and (xyz[i, : n[i]] == gp).all()
    finally:
        ctx.close()


def test_pinned_host_depth_transports_are_identical(oracle):
    """A raw 16-bit depth map in pinned host memory is either read in place over PCIe (k_unproject_host, option host_zero_copy,
    the default), uploaded whole (host_zero_copy 0), or a mix of both by call (depth_upload_of_8); pageable memory is always
    uploaded.  All must give the frames the device-resident input gives -- keypoints, descriptors and mvKeys3Dc bit for bit --
    and mvKeys3Dc must equal the oracle's back-projection."""
    import rgbdfe
    import synth
    import torch

    cam, frames = synth.sequence("tum_fr1", n=4, seed=31)
    inv = float(np.float32(1.0) / np.float32(5000.0))
    gray = [np.ascontiguousarray(f[0]) for f in frames]
    d16 = [np.clip(np.rint(f[1] * 5000.0), 0, 65535).astype(np.uint16) for f in frames]
    pin_g = [torch.from_numpy(g).pin_memory() for g in gray]
    pin_d = [torch.from_numpy(d.view(np.int16)).pin_memory() for d in d16]  # same bits as uint16
    dev_g, dev_d = [t.cuda() for t in pin_g], [t.cuda() for t in pin_d]
    ctx = rgbdfe.Context(0, 640, 480, 1000, 4)
    try:
        fr = [rgbdfe.Frame(ctx) for _ in frames]

        def run(gp, dp):
            n = len(fr)
            fa = (C.c_void_p * n)(*[f.h for f in fr])
            ctx._chk(rgbdfe.lib().rgbdfe_frames_extract_raw(ctx.h, fa, n, (C.c_void_p * n)(*gp), 640, 480, 640, (C.c_void_p * n)(*dp), 640,
                                                            C.c_float(inv), C.c_float(cam["fx"]), C.c_float(cam["fy"]), C.c_float(cam["cx"]), C.c_float(cam["cy"])))
            out = [f.download() for f in fr]
            return b"".join(a.tobytes() for o in out for a in o[:3]), out

        want, out = run([t.data_ptr() for t in dev_g], [t.data_ptr() for t in dev_d])
        orb = oracle.Orb(1000)
        for i, (g, _) in enumerate(frames):
            k, de = orb.extract(g)
            depth = d16[i].astype(np.float32) * np.float32(inv)
            assert (out[i][0] == k).all() and (out[i][1] == de).all()
            assert (out[i][2] == _unproject(cam, k, depth)).all()
        host = ([t.data_ptr() for t in pin_g], [t.data_ptr() for t in pin_d])
        for zero_copy, of8 in ((1, 0), (0, 0), (1, 4), (1, 4), (1, 8)):
            ctx.set_option("host_zero_copy", zero_copy)
            ctx.set_option("depth_upload_of_8", of8)
            ctx.set_option("depth_upload_phase", 3)  # with of8 = 4: this call uploads, the next reads in place
            for _ in range(2):
                assert run(*host)[0] == want, (zero_copy, of8)
        ctx.set_option("depth_upload_of_8", 0)
        assert run([g.ctypes.data for g in gray], [d.ctypes.data for d in d16])[0] == want  # pageable host memory
    finally:
        ctx.close()


def test_chain_replayed_as_a_graph_is_identical(oracle):
    """rgbdfe_track_sequence captures a chain's launch sequence on the second call with the same shape and replays it as one CUDA
    graph afterwards: call 1 (plain launches), call 2 (capture + launch) and call 3 (replay) must give the same bytes -- results,
    inlier lists, outlier flags, rand() position -- and the first pair must equal the oracle."""
    import rgbdfe
    import synth

    cam, frames = synth.sequence("tum_fr1", n=6, seed=777)
    ctx = rgbdfe.Context(0, 640, 480, 1000, 6)
    try:
        fr = [rgbdfe.Frame(ctx) for _ in frames]
        runs = []
        for rep in range(7):
            if rep == 3:
                ctx.set_option("chain_graphs", 0)  # and once more without graphs
            if rep == 4:  # three more with the register-capped hypothesis CTAs and no speculation (the bench's setting): plain, capture, replay
                ctx.set_option("chain_graphs", 1)
                ctx.set_option("ransac_small_ctas", 1)
                ctx.set_option("ransac_first_chunk", 1)
            ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)  # resets the outlier flags
            ctx.rng_seed(42)
            ctx.depth_cov_set(0, 0.0)
            ctx.track_sequence(fr, 0)
            res, inl = ctx.track_results_inliers(0, len(frames) - 1)
            flags = b"".join(f.download()[3].tobytes() for f in fr)
            runs.append((res.tobytes(), b"".join(x.tobytes() for x in inl), flags, ctx.rng_get_state().tobytes()))
        assert all(r == runs[0] for r in runs[1:])
        res = np.frombuffer(runs[0][0], rgbdfe.SE3_DTYPE)
        assert res["ok"].all() and res["n_inliers"].min() > 100
        orb = oracle.Orb(1000)
        k0, d0 = orb.extract(frames[0][0])
        k1, d1 = orb.extract(frames[1][0])
        p0, p1 = _unproject(cam, k0, frames[0][1]), _unproject(cam, k1, frames[1][1])
        m = oracle.match(d0, d1, np.zeros(len(k0), np.uint8), p0, p1, 0.9, True)
        r = oracle.ransac_se3(p0, p1, m, oracle.Rng(42), oracle.Latch(), 200, 10, 3.0, 4)
        assert res[0]["n_matches"] == len(m) and res[0]["n_inliers"] == len(r["inliers"]) and res[0]["iters_executed"] == r["iters_executed"]
    finally:
        ctx.close()


def test_verify_batch_matches_per_pair_oracle(oracle):
    """Config-4-shaped (small): one query vs keyframes, per-pair srand(seed + index)."""
    import rgbdfe
    import synth

    n_kf, n_feat = 24, 500
    qd, qp, kd, kp, _ = synth.synth_pair_sets(n_kf, n_feat, seed=77, overlap_frac=0.6)
    ctx = rgbdfe.Context(0, 640, 480, n_feat, 1)
    try:
        db = ctx.kfdb_create(kd, kp)
        ctx.depth_cov_set(1, (0.01 * 2.0 * 2.0) ** 2)  # latch pinned so pairs are order independent
        res = ctx.verify_batch(db, n_kf, qd, qp, 0.9, 30, 200, 30, 3.0, 4, seed=1000, first_pair_index=5)
        for k in range(n_kf):
            m = oracle.match(kd[k], qd, np.zeros(n_feat, np.uint8), kp[k], qp, 0.9, True)
            rng, latch = oracle.Rng(1000 + 5 + k), oracle.Latch()
            latch.latched.value, latch.cov.value = 1, (0.01 * 2.0 * 2.0) ** 2
            assert res[k]["n_matches"] == len(m)
            if len(m) < 30:
                assert res[k]["ok"] == 0
                continue
            r = oracle.ransac_se3(kp[k], qp, m, rng, latch, 200, 30, 3.0, 4)
            assert bool(res[k]["ok"]) == r["ok"] and res[k]["n_inliers"] == len(r["inliers"]), k
            assert res[k]["iters_executed"] == r["iters_executed"]
            assert np.abs(res[k]["T21"].reshape(4, 4) - r["T"]).max() <= POSE_TOL
        ctx.kfdb_destroy(db)
    finally:
        ctx.close()


def test_config3_full_hd_4000_features(oracle):
    """BASELINE configs[2]: 1920x1080 colour-sized frames, 4000 ORB features — extraction bit-exact, then one tracked pair."""
    import rgbdfe
    import synth

    cam, frames = synth.sequence("kinect2", n=2, seed=3456)
    ctx = rgbdfe.Context(0, 1920, 1080, 4000, 2)
    try:
        ctx.orb_configure(4000, 1.2, 8, 20, 7)
        ctx.rng_seed(42)
        fr = [rgbdfe.Frame(ctx) for _ in frames]
        ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
        ctx.track_sequence(fr, 0)
        res = ctx.track_results(0, 1)[0]
        orb = oracle.Orb(4000)
        ref = []
        for f, (g, d) in zip(fr, frames):
            k, de = orb.extract(g)
            p3 = _unproject(cam, k, d)
            gk, gd, gp, _ = f.download()
            assert len(gk) == len(k) >= 3900
            assert (gk == k).all() and (gd == de).all() and (gp == p3).all()
            ref.append((k, de, p3))
        m = oracle.match(ref[0][1], ref[1][1], np.zeros(len(ref[0][0]), np.uint8), ref[0][2], ref[1][2], 0.9, True)
        r = oracle.ransac_se3(ref[0][2], ref[1][2], m, oracle.Rng(42), oracle.Latch(), 200, 10, 3.0, 4)
        assert res["n_matches"] == len(m) and bool(res["ok"]) == r["ok"] and res["n_inliers"] == len(r["inliers"])
        assert res["iters_executed"] == r["iters_executed"]
        assert np.abs(res["T21"].reshape(4, 4) - r["T"]).max() <= POSE_TOL
    finally:
        ctx.close()


def test_config4_full_size_shard_independence_and_poses(oracle):
    """BASELINE configs[3] at full size: 1 query x 4096 keyframes x 1000 features.  Size-independent properties:
    the result of a keyframe does not depend on how the database is sharded (whole database == two halves with
    first_pair_index offsets, byte for byte); overlapping keyframes verify with a pose that inverts the generator's
    transform; a sample of pairs equals the oracle."""
    import rgbdfe
    import synth

    n_kf, n_feat = 4096, 1000
    qd, qp, kd, kp, T = synth.synth_pair_sets(n_kf, n_feat, seed=5678, overlap_frac=0.05)
    cov = (0.01 * 2.0 * 2.0) ** 2
    ctx = rgbdfe.Context(0, 640, 480, n_feat, 1)
    try:
        ctx.depth_cov_set(1, cov)
        db = ctx.kfdb_create(kd, kp)
        whole = ctx.verify_batch(db, n_kf, qd, qp, 0.9, 30, 200, 30, 3.0, 4, seed=42, first_pair_index=0)
        ctx.kfdb_destroy(db)
        halves = []
        for lo, hi in ((0, 2048), (2048, 4096)):
            db = ctx.kfdb_create(kd[lo:hi], kp[lo:hi])
            halves.append(ctx.verify_batch(db, hi - lo, qd, qp, 0.9, 30, 200, 30, 3.0, 4, seed=42, first_pair_index=lo))
            ctx.kfdb_destroy(db)
        assert np.concatenate(halves).tobytes() == whole.tobytes()
        overlap = np.nonzero(np.abs(T - np.eye(4, dtype=np.float32)).reshape(n_kf, -1).max(1) > 0)[0]
        assert 150 < len(overlap) < 260
        assert whole["ok"][overlap].mean() > 0.95 and whole["ok"].sum() <= len(overlap)  # no false loop closures
        for k in overlap[whole["ok"][overlap] != 0][:64]:
            err = whole[k]["T21"].reshape(4, 4).astype(np.float64) @ T[k].astype(np.float64) - np.eye(4)
            assert np.abs(err).max() < 0.05, k  # sensor noise 0.0012 z^2 on ~400 correspondences
        for k in list(overlap[:6]) + [0, 1, 2]:
            m = oracle.match(kd[k], qd, np.zeros(n_feat, np.uint8), kp[k], qp, 0.9, True)
            assert whole[k]["n_matches"] == len(m)
            if len(m) < 30:
                assert whole[k]["ok"] == 0
                continue
            rng, latch = oracle.Rng(42 + int(k)), oracle.Latch()
            latch.latched.value, latch.cov.value = 1, cov
            r = oracle.ransac_se3(kp[k], qp, m, rng, latch, 200, 30, 3.0, 4)
            assert bool(whole[k]["ok"]) == r["ok"] and whole[k]["n_inliers"] == len(r["inliers"]), k
            assert np.abs(whole[k]["T21"].reshape(4, 4) - r["T"]).max() <= POSE_TOL
    finally:
        ctx.close()


def test_config5_batched_relocalisation_queries_are_independent(oracle):
    """BASELINE configs[4] shape (Q queries x 1024 keyframes, all pairs) at full keyframe count and Q = 12: running the
    queries one after the other on one context gives, for every query, exactly what a fresh context gives for it
    alone (per-pair streams srand(seed + q * n_kf + k): no state leaks from query to query)."""
    import rgbdfe
    import synth

    n_kf, n_feat, Q = 1024, 1000, 12
    qd, qp, kd, kp, _ = synth.synth_pair_sets(n_kf, n_feat, seed=6789, overlap_frac=0.1)
    rs = np.random.default_rng(1)
    queries = [(qd, qp)] + [(lambda p: (qd[p], qp[p]))(rs.permutation(n_feat)) for _ in range(Q - 1)]
    cov = (0.01 * 2.0 * 2.0) ** 2

    def run(ctx, db, q):
        return ctx.verify_batch(db, n_kf, queries[q][0], queries[q][1], 0.9, 30, 200, 30, 3.0, 4, seed=7, first_pair_index=q * n_kf)

    ctx = rgbdfe.Context(0, 640, 480, n_feat, 1)
    ctx.depth_cov_set(1, cov)
    db = ctx.kfdb_create(kd, kp)
    batch = [run(ctx, db, q) for q in range(Q)]
    ctx.kfdb_destroy(db)
    ctx.close()
    assert all(b["ok"].sum() > 50 for b in batch)
    for q in (0, 5, Q - 1):
        c2 = rgbdfe.Context(0, 640, 480, n_feat, 1)
        c2.depth_cov_set(1, cov)
        d2 = c2.kfdb_create(kd, kp)
        alone = run(c2, d2, q)
        c2.kfdb_destroy(d2)
        c2.close()
        assert alone.tobytes() == batch[q].tobytes()
    # a permuted query sees the same scene: the same keyframes close their loops
    assert (batch[3]["ok"] != 0).sum() == (batch[0]["ok"] != 0).sum() or abs(int(batch[3]["ok"].sum()) - int(batch[0]["ok"].sum())) <= 3
