"""Golden vectors (tests/golden/, made by tests/golden/make_golden.py).

cv2_primitives.npz  : outputs of the real OpenCV 4.13 / glibc / libstdc++ code the reference calls on the path.
pipeline_oracle.npz : oracle outputs for a small tracked sequence (regression vectors, the reference itself does not
                      build in this image).
CPU tests pin the oracle to both files; GPU tests run the CUDA path (through the C ABI) against the same files, so
the GPU parity does not rest on the oracle alone.
"""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def prim():
    return np.load(os.path.join(GOLD, "cv2_primitives.npz"))


@pytest.fixture(scope="module")
def pipe():
    return np.load(os.path.join(GOLD, "pipeline_oracle.npz"))


def _filter_from_knn(idx, dist, ratio):
    """Matcher::match's filter (Features/Matcher.cpp:114-136) on kNN results, all observations valid, no outliers."""
    used, out = set(), []
    for q in range(len(idx)):
        if np.float32(dist[q, 0]) < np.float32(ratio) * np.float32(dist[q, 1]):
            t = int(idx[q, 0])
            if t in used:
                continue
            used.add(t)
            out.append((q, t, dist[q, 0]))
    return out


# ------------------------------------------------------------------ CPU: oracle vs golden
def test_oracle_image_primitives_vs_cv2_golden(oracle, prim):
    sizes = prim["level_sizes"]
    assert (sizes[0] == [320, 240]).all()
    for l in range(1, 8):
        w, h = (int(v) for v in sizes[l])
        assert (oracle.resize_linear(prim[f"level{l - 1}"], w, h) == prim[f"level{l}"]).all(), f"resize level {l}"
    for l in range(8):
        assert (oracle.gauss7(prim[f"level{l}"]) == prim[f"blur{l}"]).all(), f"blur level {l}"
    assert (oracle.border101(prim["level3"], 19) == prim["border19_level3"]).all()
    for k in range(3):
        for th in (7, 20):
            for nm in (0, 1):
                ref = prim[f"fast{k}_th{th}_nm{nm}"]
                got = oracle.fast9(prim[f"fast_img{k}"], th, bool(nm)).astype(np.float32).reshape(-1, 3)
                assert ref.shape == got.shape and (ref == got).all(), f"FAST image {k} th {th} nonmax {nm}"
    assert (oracle.fast_atan2_v(prim["atan2_y"], prim["atan2_x"]) == prim["atan2_deg"]).all()


def test_oracle_knn_rand_sort_vs_golden(oracle, prim):
    idx, dist = oracle.knn2(prim["knn_q"], prim["knn_t"])
    assert (idx == prim["knn_idx"]).all() and (dist == prim["knn_dist"]).all()
    for seed in (42, 1, 0, 123456789):
        r = oracle.Rng(seed)
        ref = prim[f"rand_seed{seed}"]
        assert all(int(v) == r.next() for v in ref), f"rand() stream of seed {seed}"
    import orclib

    for k in range(int(prim["sort_cases"])):
        keys, perm = prim[f"sort_keys{k}"], prim[f"sort_perm{k}"]
        m = np.zeros(len(keys), orclib.DM_DTYPE)
        m["distance"], m["queryIdx"] = keys, np.arange(len(keys))
        assert (oracle.sort_matches(m)["queryIdx"] == perm).all(), f"std::sort permutation case {k}"


def _pipe_frames(pipe):
    cam = dict(zip(("fx", "fy", "cx", "cy", "factor"), (float(v) for v in pipe["cam"])))
    frames = []
    for i in range(3):
        depth = pipe[f"depth16_{i}"].astype(np.float32) * np.float32(1.0 / 5000.0)
        frames.append((pipe[f"gray{i}"], depth))
    return cam, frames


def test_oracle_pipeline_regression(oracle, pipe):
    import synth

    cam, frames = _pipe_frames(pipe)
    orb = oracle.Orb(int(pipe["nfeatures"]))
    rng, latch = oracle.Rng(42), oracle.Latch()
    prev = None
    for i, (g, d) in enumerate(frames):
        k, de = orb.extract(g)
        for f in ("x", "y", "size", "response", "octave"):
            assert (k[f] == pipe[f"kp{i}_{f}"]).all(), f"frame {i} keypoint {f}"
        assert (k["angle"].view(np.uint32) == pipe[f"kp{i}_angle"].view(np.uint32)).all()
        assert (de == pipe[f"desc{i}"]).all()
        p3 = synth.unproject(cam, np.stack([k["x"], k["y"]], 1), d)
        assert (p3 == pipe[f"xyz{i}"]).all()
        if prev is not None:
            m = oracle.match(prev[1], de, prev[3], prev[2], p3, 0.9, True)
            assert (m["queryIdx"] == pipe[f"match{i}_q"]).all() and (m["trainIdx"] == pipe[f"match{i}_t"]).all()
            r = oracle.ransac_se3(prev[2], p3, m, rng, latch, 200, 10, 3.0, 4)
            assert int(r["ok"]) == int(pipe[f"se3_{i}_ok"]) and r["iters_executed"] == int(pipe[f"se3_{i}_iters"])
            assert (r["inliers"]["trainIdx"] == pipe[f"se3_{i}_inl_t"]).all()
            assert (r["T"] == pipe[f"se3_{i}_T"]).all()
        prev = (k, de, p3, pipe[f"outlier{i}"])
    assert rng.next() == int(pipe["rng_next_after"])


# ------------------------------------------------------------------ GPU: CUDA path vs golden
@pytest.mark.gpu
def test_cuda_pyramid_blur_knn_rand_vs_cv2_golden(prim):
    import rgbdfe

    ctx = rgbdfe.Context(0, 320, 240, 500, 1)
    try:
        ctx.orb_configure(500, 1.2, 8, 20, 7)
        ctx.orb_extract(prim["img"])
        for l in range(8):
            assert (ctx.pyramid_level(l)[19:-19, 19:-19] == prim[f"level{l}"]).all(), f"pyramid level {l} vs cv::resize chain"
            assert (ctx.stage_blurred(l) == prim[f"blur{l}"]).all(), f"blur level {l} vs cv::GaussianBlur"
        assert (ctx.pyramid_level(3) == prim["border19_level3"]).all()
        q, t = prim["knn_q"], prim["knn_t"]
        ones_q, ones_t = np.ones((len(q), 3), np.float32), np.ones((len(t), 3), np.float32)
        for ratio in (0.9, 0.7):
            ref = _filter_from_knn(prim["knn_idx"], prim["knn_dist"], ratio)
            got = ctx.match(q, t, None, ones_q, ones_t, ratio, False)
            assert [(int(m["queryIdx"]), int(m["trainIdx"]), float(m["distance"])) for m in got] == [(a, b, float(c)) for a, b, c in ref]
        for seed in (42, 0, 123456789):
            ctx.rng_seed(seed)
            ref = prim[f"rand_seed{seed}"][:300]
            assert [ctx.rng_next() for _ in range(len(ref))] == [int(v) for v in ref]
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("raw_depth", [False, True, "in_place"])
def test_cuda_pipeline_vs_golden(pipe, raw_depth):
    """raw_depth: the 16-bit depth image goes to the device and Frame's convertTo(CV_32F, 1/factor) happens there (N1);
    in_place: the same with images that already live in device memory (read where they are, no staging copy)."""
    import ctypes as C

    import rgbdfe

    cam, frames = _pipe_frames(pipe)
    ctx = rgbdfe.Context(0, 320, 240, int(pipe["nfeatures"]), 4)
    try:
        ctx.orb_configure(int(pipe["nfeatures"]), 1.2, 8, 20, 7)
        ctx.rng_seed(42)
        fr = [rgbdfe.Frame(ctx) for _ in frames]
        inv = float(np.float32(1.0) / np.float32(cam["factor"]))
        if raw_depth == "in_place":
            import torch

            g_d = [torch.from_numpy(np.ascontiguousarray(f[0])).cuda() for f in frames]
            d_d = [torch.from_numpy(pipe[f"depth16_{i}"].astype(np.int16)).cuda() for i in range(3)]  # same bits as uint16
            torch.cuda.synchronize()
            n = len(frames)
            ctx._chk(rgbdfe.lib().rgbdfe_frames_extract_raw(
                ctx.h, (C.c_void_p * n)(*[f.h for f in fr]), n, (C.c_void_p * n)(*[t.data_ptr() for t in g_d]), 320, 240, 320,
                (C.c_void_p * n)(*[t.data_ptr() for t in d_d]), 320, C.c_float(inv), C.c_float(cam["fx"]), C.c_float(cam["fy"]),
                C.c_float(cam["cx"]), C.c_float(cam["cy"])))
        elif raw_depth:
            d16 = [np.ascontiguousarray(pipe[f"depth16_{i}"]) for i in range(3)]
            ctx.frames_extract_raw(fr, [f[0] for f in frames], d16, cam, inv)
        else:
            ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
        ctx.track_sequence(fr, 0)
        res = ctx.track_results(0, 2)
        for i in range(3):
            k, de, p3, outl = fr[i].download()
            for f in ("x", "y", "size", "response", "octave"):
                assert (k[f] == pipe[f"kp{i}_{f}"]).all(), f"frame {i} keypoint {f}"
            assert (k["angle"].view(np.uint32) == pipe[f"kp{i}_angle"].view(np.uint32)).all()
            assert (de == pipe[f"desc{i}"]).all() and (p3 == pipe[f"xyz{i}"]).all()
            assert (outl == pipe[f"outlier{i}"]).all()
            if i:
                r = res[i - 1]
                assert r["n_matches"] == len(pipe[f"match{i}_q"]) and int(r["ok"]) == int(pipe[f"se3_{i}_ok"])
                assert r["n_inliers"] == len(pipe[f"se3_{i}_inl_t"]) and r["iters_executed"] == int(pipe[f"se3_{i}_iters"])
                assert np.abs(r["T21"].reshape(4, 4) - pipe[f"se3_{i}_T"]).max() <= 1e-5
        assert ctx.rng_next() == int(pipe["rng_next_after"])
    finally:
        ctx.close()
