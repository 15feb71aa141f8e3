"""Row N3: Extractor(ORB, ORB) = OpenCV's cv::ORB with Harris scoring
(Features/Extractor.cpp:55-59,163-166,216-218).

CPU: the restatement oracle/cvorb.cpp against cv2 4.13 -- live (cv2.ORB_create(...).detect + cv2.ORB_create().compute,
resize(INTER_LINEAR_EXACT), the float sepFilter2D blur) and through tests/golden/cvorb_cv2.npz.
GPU: rgbdfe_cvorb_configure + the extraction entry points (C ABI) against the golden cv2 vectors and the oracle,
stage by stage and end to end, bit for bit (keypoint order included: it is libstdc++'s nth_element permutation).
"""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")]
GOLD = os.path.join(ROOT, "tests", "golden")

import orclib  # noqa: E402
import synth  # noqa: E402


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "cvorb_cv2.npz"))


def _same_kps(a, b):
    return len(a) == len(b) and all((a[f].view(np.uint32) == b[f].view(np.uint32)).all() for f in a.dtype.names)


def _blur_cols(w):
    """columns of the blurred level that are pinned: all but the last w % 4 (sepFilter2D's scalar tail, never sampled)"""
    return slice(0, w - w % 4)


# ------------------------------------------------------------------ CPU: oracle vs OpenCV
def test_introselect_restatement_matches_libstdcxx(tmp_path):
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "cpp", "test_introselect.cpp")])
    subprocess.check_call([str(exe)])


def test_oracle_primitives_vs_cv2_golden(gold):
    assert (orclib.resize_linear_exact(gold["resize_src"], 109, 81) == gold["resize_dst"]).all()
    w = gold["blur_src"].shape[1]
    assert (orclib.gauss7_float(gold["blur_src"])[:, _blur_cols(w)] == gold["blur_dst"][:, _blur_cols(w)]).all()


@pytest.mark.parametrize("case,nf", [("a", 500), ("b", 300)])
def test_oracle_cvorb_vs_cv2_golden(gold, case, nf):
    k, d = orclib.CvOrb(nf, 1.2, 8).extract(gold["img_" + case])
    assert _same_kps(k, gold["kps_" + case])
    assert (d == gold["desc_" + case]).all()


def test_oracle_cvorb_golden_b_has_ties(gold):
    """the periodic fixture makes both retainBest calls cut through equal responses, so the order is really exercised"""
    k = gold["kps_b"]
    l0 = k[k["octave"] == 0]
    assert len(np.unique(l0["response"])) < len(l0) // 2


def test_oracle_vs_cv2_live():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for _ in range(8):
        sh, sw = int(rng.integers(8, 300)), int(rng.integers(8, 400))
        dh, dw = int(rng.integers(4, sh + 1)), int(rng.integers(4, sw + 1))
        src = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        assert (cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR_EXACT) == orclib.resize_linear_exact(src, dw, dh)).all()
    cam, frames = synth.sequence("icl", n=1, seed=31)
    g = frames[0][0]
    det, des = cv2.ORB_create(1000, 1.2, 8), cv2.ORB_create()
    k = det.detect(g, None)
    k, d = des.compute(g, k)
    ck = np.array([(p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave, p.class_id) for p in k], dtype=orclib.KP_DTYPE)
    ok, od = orclib.CvOrb(1000, 1.2, 8).extract(g)
    assert _same_kps(ok, ck) and (od == d).all()


# ------------------------------------------------------------------ GPU: CUDA path vs golden / oracle
def _ctx(w, h, nf, batch=1):
    import rgbdfe

    ctx = rgbdfe.Context(0, w, h, max(nf, 1000), batch)
    ctx.cvorb_configure(nf, 1.2, 8)
    return ctx


@pytest.mark.gpu
@pytest.mark.parametrize("case,nf", [("a", 500), ("b", 300)])
def test_gpu_cvorb_vs_cv2_golden(gold, case, nf):
    img = gold["img_" + case]
    ctx = _ctx(img.shape[1], img.shape[0], nf)
    k, d = ctx.orb_extract(img)
    assert _same_kps(k, gold["kps_" + case])
    assert (d == gold["desc_" + case]).all()
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,nf,seed", [("tum_fr1", 1000, 11), ("icl", 2000, 12)])
def test_gpu_cvorb_stages_vs_oracle(name, nf, seed):
    cam, frames = synth.sequence(name, n=1, seed=seed)
    g = frames[0][0]
    o = orclib.CvOrb(nf, 1.2, 8)
    ok, od = o.extract(g)
    ctx = _ctx(640, 480, nf)
    k, d = ctx.orb_extract(g)
    for l in range(8):
        info = o.level_info(l)
        assert (ctx.cvorb_stage_level(l, 0) == o.level_image(l)).all(), f"pyramid level {l}"
        assert (ctx.cvorb_stage_level(l, 1)[:, _blur_cols(info["w"])] == o.level_image(l, True)[:, _blur_cols(info["w"])]).all(), f"blur {l}"
        assert int((ctx.cvorb_stage_level(l, 2) != 0).sum()) == info["n_fast"], f"FAST count {l}"
        hg, ho = ctx.cvorb_stage_harris(l), o.level_harris(l)
        assert len(hg) == len(ho) and (hg["x"] == ho["x"]).all() and (hg["y"] == ho["y"]).all(), f"retainBest(2N) order {l}"
        assert (hg["response"].view(np.uint32) == ho["response"].view(np.uint32)).all(), f"Harris response {l}"
    assert _same_kps(k, ok)
    assert (d == od).all()
    ctx.close()


@pytest.mark.gpu
def test_gpu_cvorb_batched_frames_and_tracking():
    """the cv::ORB branch behind the batched frame entry point: keypoints, descriptors, back-projected points, then
    Matcher::match + RansacSE3 on those frames against the oracle fed with the oracle's cv::ORB features"""
    import rgbdfe

    cam, frames = synth.sequence("tum_fr1", n=3, seed=41)
    ctx = _ctx(640, 480, 1000, batch=3)
    ctx.rng_seed(42)
    fr = [rgbdfe.Frame(ctx) for _ in frames]
    ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
    o = orclib.CvOrb(1000, 1.2, 8)
    ref = []
    for (g, dep), f in zip(frames, fr):
        k, de = o.extract(g)
        p3 = synth.unproject(cam, np.stack([k["x"], k["y"]], 1), dep)
        gk, gd, gp, _ = f.download()
        assert _same_kps(gk, k) and (gd == de).all() and (gp == p3).all()
        ref.append((k, de, p3))
    rng, latch = orclib.Rng(42), orclib.Latch()
    ctx.track(fr[0], fr[1], 0)
    res = ctx.track_results(0, 1)[0]
    m = orclib.match(ref[0][1], ref[1][1], np.zeros(len(ref[0][0]), np.uint8), ref[0][2], ref[1][2], 0.9, True)
    r = orclib.ransac_se3(ref[0][2], ref[1][2], m, rng, latch)
    assert res["n_matches"] == len(m) and bool(res["ok"]) == r["ok"] and res["n_inliers"] == len(r["inliers"])
    assert np.abs(res["T21"].reshape(4, 4) - r["T"]).max() <= 1e-5  # north_star tolerance (FP32 pose)
    for f in fr:
        f.close()
    ctx.close()


@pytest.mark.gpu
def test_gpu_cvorb_1080p_and_switch_back():
    """config 3 size (1920x1080, 4000 features) and switching the context back to the ORB2 extractor"""
    cam, frames = synth.sequence("kinect2", n=1, seed=51)
    g = frames[0][0]
    ctx = _ctx(g.shape[1], g.shape[0], 4000)
    k, d = ctx.orb_extract(g)
    ok, od = orclib.CvOrb(4000, 1.2, 8).extract(g)
    assert _same_kps(k, ok) and (d == od).all()
    ctx.orb_configure(1000, 1.2, 8, 20, 7)
    k2, d2 = ctx.orb_extract(g)
    o2k, o2d = orclib.Orb(1000).extract(g)
    assert _same_kps(k2, o2k) and (d2 == o2d).all()
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nf", [(517, 333, 500), (322, 241, 40), (640, 480, 3)])
def test_gpu_cvorb_odd_sizes_and_tiny_quotas(w, h, nf):
    """widths / heights that are not multiples of 4 (row pitch padding, word staging) and feature counts so small that
    some levels get a quota of 0 (retainBest(0) clears the level) or 1"""
    cv2 = pytest.importorskip("cv2")
    cam, frames = synth.sequence("tum_fr1", n=1, seed=61)
    g = np.ascontiguousarray(cv2.resize(frames[0][0], (w, h), interpolation=cv2.INTER_AREA))
    ok, od = orclib.CvOrb(nf, 1.2, 8).extract(g)
    ctx = _ctx(w, h, nf)
    k, d = ctx.orb_extract(g)
    assert _same_kps(k, ok) and (d == od).all()
    # a strided view (row pitch != width) gives the same result
    big = np.zeros((h, w + 13), np.uint8)
    big[:, :w] = g
    import ctypes as C

    import rgbdfe

    kps = np.zeros(ctx.kp_cap, rgbdfe.KP_DTYPE)
    desc = np.zeros((ctx.kp_cap, 32), np.uint8)
    n = C.c_int()
    ctx._chk(rgbdfe.lib().rgbdfe_orb_extract(ctx.h, big.ctypes.data_as(C.c_void_p), w, h, big.strides[0], kps.ctypes.data_as(C.c_void_p),
                                             desc.ctypes.data_as(C.c_void_p), ctx.kp_cap, C.byref(n)))
    assert _same_kps(kps[: n.value], ok) and (desc[: n.value] == od).all()
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("blurred", [True, False])
def test_gpu_cvorb_many_corners_takes_the_global_memory_path(blurred):
    """a noisy image yields more FAST corners per level than the selection kernel keeps in shared memory (8192 at
    640x480): the same permutation-exact retainBest then runs on the global scratch"""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    g = rng.integers(0, 256, (480, 640), dtype=np.uint8)
    if blurred:
        g = cv2.GaussianBlur(g, (3, 3), 0.8)
    o = orclib.CvOrb(1000, 1.2, 8)
    ok, od = o.extract(g)
    assert o.level_info(0)["n_fast"] > 8192, o.level_info(0)
    ctx = _ctx(640, 480, 1000)
    k, d = ctx.orb_extract(g)
    assert _same_kps(k, ok) and (d == od).all()
    ctx.close()


@pytest.mark.gpu
def test_gpu_cvorb_capacity_overflow_is_an_error_not_garbage():
    """white noise at 1920x1080 has > 65535 FAST corners on level 0 (16-bit stopper positions): the call must fail"""
    import rgbdfe

    g = np.random.default_rng(6).integers(0, 256, (1080, 1920), dtype=np.uint8)
    ctx = _ctx(1920, 1080, 1000)
    with pytest.raises(rgbdfe.RgbdfeError):
        ctx.orb_extract(g)
    ctx.close()
