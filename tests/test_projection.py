"""Row N2: Matcher::projectionMatch (Features/Matcher.cpp:35-104, Frame::getFeaturesInArea Core/Frame.cpp:179-229).

CPU: the numpy oracle (oracle/projection.py) -- its cv::Mat arithmetic against cv2.gemm / cv2.norm, live and through
     tests/golden/projection_cv2.npz; behaviour of the restated function on constructed scenes.
GPU: rgbdfe_projection_match (host arrays) and rgbdfe_frames_projection_match (device-resident frames) through the
     C ABI against the oracle, match for match.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")

CAM = dict(fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989)
W, H = 640, 480


def _rot(rng, s):
    """Small rotation (Rodrigues) as float32."""
    w = rng.normal(0, s, 3)
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.eye(3, dtype=np.float32)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return (np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K).astype(np.float32)


def make_scene(seed, n1=700, n2=800, motion=0.02, flips=0.06, pool=0, invalid=0.1, outliers=0.1, bounds=None, inv=None):
    """Two frames that see the same random 3-D points.  pool > 0: descriptors come from a pool of that many distinct
    rows, so equal distances are common and the reference's traversal order decides."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    bounds = bounds or (f32(0), f32(W), f32(0), f32(H))
    inv = inv or (f32(64) / f32(bounds[1] - bounds[0]), f32(48) / f32(bounds[3] - bounds[2]))
    # frame 1: world pose Tcw1, points in its camera frame
    R1, t1 = _rot(rng, 0.05), rng.normal(0, 0.1, 3).astype(f32)
    uv1 = np.stack([rng.uniform(5, W - 5, n1), rng.uniform(5, H - 5, n1)], 1).astype(f32)
    z1 = rng.uniform(0.6, 5.0, n1).astype(f32)
    xyz1 = np.stack([(uv1[:, 0] - f32(CAM["cx"])) * z1 / f32(CAM["fx"]), (uv1[:, 1] - f32(CAM["cy"])) * z1 / f32(CAM["fy"]), z1], 1).astype(f32)
    xyz1[rng.random(n1) < invalid] = 0
    Rwc1 = np.ascontiguousarray(R1.T)
    Ow1 = (-(R1.T.astype(np.float64) @ t1.astype(np.float64))).astype(f32)
    # frame 2: a little further along
    R2 = (_rot(rng, motion).astype(np.float64) @ R1.astype(np.float64)).astype(f32)
    t2 = (t1 + rng.normal(0, motion, 3)).astype(f32)
    world = (Rwc1.astype(np.float64) @ xyz1.T.astype(np.float64)).T + Ow1
    cam2 = (R2.astype(np.float64) @ world.T).T + t2
    with np.errstate(all="ignore"):
        uv2 = np.stack([CAM["fx"] * cam2[:, 0] / cam2[:, 2] + CAM["cx"], CAM["fy"] * cam2[:, 1] / cam2[:, 2] + CAM["cy"]], 1)
    if pool:
        bank = rng.integers(0, 256, (pool, 32), dtype=np.uint8)
        desc1 = bank[rng.integers(0, pool, n1)]
    else:
        desc1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    # frame 2 keypoints: noisy re-observations of frame 1's points (shuffled) plus unrelated ones
    src = rng.permutation(n1)[: min(n1, n2 * 3 // 4)]
    k = len(src)
    pts2 = np.empty((n2, 2), f32)
    desc2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8) if not pool else bank[rng.integers(0, pool, n2)]
    pts2[:k] = np.nan_to_num(uv2[src]) + rng.normal(0, 1.5, (k, 2))
    flip = np.packbits(rng.random((k, 256)) < flips, axis=1)
    desc2[:k] = desc1[src] ^ flip
    pts2[k:] = np.stack([rng.uniform(-10, W + 10, n2 - k), rng.uniform(-10, H + 10, n2 - k)], 1)
    order = rng.permutation(n2)
    pts2, desc2 = np.ascontiguousarray(pts2[order]), np.ascontiguousarray(desc2[order])
    z2 = rng.uniform(0.6, 5.0, n2).astype(f32)
    xyz2 = np.stack([pts2[:, 0] * z2, pts2[:, 1] * z2, z2], 1).astype(f32)  # only z is read for F2
    xyz2[rng.random(n2) < invalid] = 0
    f1 = dict(desc=np.ascontiguousarray(desc1), xyz=xyz1, outlier=(rng.random(n1) < outliers))
    f2 = dict(desc=desc2, xyz=xyz2, pts_un=pts2)
    proj = dict(CAM, Rwc1=Rwc1, Ow1=Ow1, Rcw2=R2, tcw2=t2, bounds=bounds, inv=inv)
    return f1, f2, proj


# ------------------------------------------------------------------ CPU
def test_oracle_arithmetic_vs_cv2_golden():
    import projection as pj

    g = np.load(os.path.join(GOLD, "projection_cv2.npz"))
    got = np.stack([pj.mat_vec_add(g["A"][i], g["b"][i], g["c"][i]) for i in range(len(g["A"]))])
    assert (got.view(np.uint32) == g["gemm"].view(np.uint32)).all()
    assert [pj.hamming(a, b) for a, b in zip(g["desc_a"], g["desc_b"])] == [int(v) for v in g["hamming"]]


def test_oracle_arithmetic_vs_cv2_live():
    cv2 = pytest.importorskip("cv2")
    import projection as pj

    rng = np.random.default_rng(5)
    for _ in range(500):
        A = rng.normal(0, 1, (3, 3)).astype(np.float32)
        b = rng.normal(0, 3, (3, 1)).astype(np.float32)
        c = rng.normal(0, 1, (3, 1)).astype(np.float32)
        ref = cv2.gemm(A, b, 1.0, c, 1.0).reshape(3)
        assert (pj.mat_vec_add(A, b, c).view(np.uint32) == ref.view(np.uint32)).all()


def test_oracle_projection_match_behaviour():
    import projection as pj

    f1, f2, proj = make_scene(1, n1=300, n2=320)
    m = pj.projection_match(f1, f2, proj, 5.0)
    assert len(m) > 50
    q, t, d = (np.array(v) for v in zip(*m))
    assert (np.diff(q) > 0).all(), "query order"
    assert len(set(t.tolist())) == len(t), "a train keypoint is taken once"
    assert (d <= pj.TH_HIGH).all() and not f1["outlier"][q].any() and (f1["xyz"][q, 2] > 0).all() and (f2["xyz"][t, 2] > 0).all()
    assert [pj.hamming(f1["desc"][a], f2["desc"][b]) for a, b in zip(q, t)] == d.astype(int).tolist()
    # same frame, same pose: every valid inlier keypoint inside the image finds itself at distance 0
    same = dict(desc=f1["desc"], xyz=f1["xyz"], pts_un=_project(f1["xyz"]))
    p2 = dict(proj, Rcw2=proj["Rwc1"].T.copy(), tcw2=(-(proj["Rwc1"].T.astype(np.float64) @ proj["Ow1"].astype(np.float64))).astype(np.float32))
    m2 = pj.projection_match(f1, same, p2, 5.0)
    ok = (~f1["outlier"]) & (f1["xyz"][:, 2] > 0)
    assert [a for a, _, _ in m2] == np.nonzero(ok)[0].tolist() and all(a == b and d == 0 for a, b, d in m2)
    # radius 0 admits nobody (|dx| < r is strict)
    assert pj.projection_match(f1, f2, proj, 0.0) == []


def _project(xyz):
    with np.errstate(all="ignore"):
        u = np.float32(CAM["fx"]) * xyz[:, 0] / xyz[:, 2] + np.float32(CAM["cx"])
        v = np.float32(CAM["fy"]) * xyz[:, 1] / xyz[:, 2] + np.float32(CAM["cy"])
    return np.nan_to_num(np.stack([u, v], 1)).astype(np.float32)


def test_oracle_features_in_area_cell_range():
    """getFeaturesInArea keeps the quirks of the reference: round() for the cell of a keypoint, floor()/ceil() for the
    range searched, so a keypoint just inside the window can sit in a cell outside the range and is then not found."""
    import projection as pj

    bounds, inv = (0.0, 640.0, 0.0, 480.0), (np.float32(0.1), np.float32(0.1))
    pts = np.array([[100.0, 100.0], [104.9, 100.0], [95.1, 100.0], [100.0, 109.0]], np.float32)
    grid = pj.build_grid(pts, bounds, inv)
    assert grid == {(10, 10): [0, 1, 2], (10, 11): [3]}
    assert pj.features_in_area(grid, pts, bounds, inv, 100.0, 100.0, 5.0) == [0, 1, 2]
    assert pj.features_in_area(grid, pts, bounds, inv, 100.0, 100.0, 9.5) == [0, 1, 2, 3]
    assert pj.features_in_area(grid, pts, bounds, inv, -50.0, 100.0, 5.0) == []
    assert pj.features_in_area(grid, pts, bounds, inv, float("nan"), 100.0, 5.0) == []


# ------------------------------------------------------------------ GPU
def _as_tuples(m):
    return [(int(a["queryIdx"]), int(a["trainIdx"]), float(a["distance"])) for a in m]


CASES = {
    "default_radius": dict(seed=11, th=5.0),
    "wide_radius": dict(seed=12, th=25.0),
    "descriptor_pool_ties": dict(seed=13, th=30.0, pool=6, flips=0.0, n1=900, n2=900),
    "crowded_claims": dict(seed=14, th=80.0, pool=3, flips=0.01, n1=500, n2=260),
    "whole_image": dict(seed=15, th=1.0e4, n1=200, n2=220, flips=0.15),
    "absurd_radius": dict(seed=16, th=3.0e10, n1=120, n2=100),
    "tiny": dict(seed=17, th=5.0, n1=3, n2=2),
    "undistorted_bounds": dict(seed=18, th=7.5, bounds=(-28.4, 670.2, -25.1, 511.3)),
}


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_cuda_projection_match_host_arrays(name):
    import projection as pj
    import rgbdfe

    kw = dict(CASES[name])
    th = kw.pop("th")
    if "bounds" in kw:
        kw["bounds"] = tuple(np.float32(v) for v in kw["bounds"])
    f1, f2, proj = make_scene(**kw)
    ref = pj.projection_match(f1, f2, proj, th)
    ctx = rgbdfe.Context(0, W, H, 1000, 1)
    try:
        got = ctx.projection_match(f1["desc"], f1["xyz"], f1["outlier"], f2["desc"], f2["xyz"], f2["pts_un"], proj, th)
        assert _as_tuples(got) == ref, name
        assert (got["imgIdx"] == -1).all()
        if name in ("descriptor_pool_ties", "crowded_claims"):
            assert ctx.projection_rounds(1)[0] > 2, "the scene is meant to make later keypoints lose their first choice"
        # no outlier flags given: every keypoint of F1 is an inlier
        ref2 = pj.projection_match(dict(f1, outlier=None), f2, proj, th)
        got2 = ctx.projection_match(f1["desc"], f1["xyz"], None, f2["desc"], f2["xyz"], f2["pts_un"], proj, th)
        assert _as_tuples(got2) == ref2
    finally:
        ctx.close()


@pytest.mark.gpu
def test_cuda_projection_match_edges():
    import projection as pj
    import rgbdfe

    f1, f2, proj = make_scene(21, n1=150, n2=160)
    ctx = rgbdfe.Context(0, W, H, 1000, 1)
    try:
        run = lambda a, b, p, th: _as_tuples(ctx.projection_match(a["desc"], a["xyz"], a["outlier"], b["desc"], b["xyz"], b["pts_un"], p, th))
        assert run(dict(f1, desc=f1["desc"][:0], xyz=f1["xyz"][:0], outlier=f1["outlier"][:0]), f2, proj, 5.0) == []
        assert run(f1, f2, proj, 0.0) == [] and run(f1, f2, proj, -3.0) == []
        # points behind F2 (invzc < 0), on its focal plane (division by zero) and NaN poses
        behind = dict(proj, tcw2=(proj["tcw2"] - np.float32([0, 0, 50])).astype(np.float32))
        assert run(f1, f2, behind, 5.0) == pj.projection_match(f1, f2, behind, 5.0) == []
        flat = dict(proj, Rcw2=np.zeros((3, 3), np.float32), tcw2=np.zeros(3, np.float32))
        assert run(f1, f2, flat, 5.0) == pj.projection_match(f1, f2, flat, 5.0) == []
        nanp = dict(proj, tcw2=np.float32([np.nan, 0, 1]))
        assert run(f1, f2, nanp, 5.0) == pj.projection_match(f1, f2, nanp, 5.0)
        # every F2 observation invalid
        dead = dict(f2, xyz=np.zeros_like(f2["xyz"]))
        assert run(f1, dead, proj, 5.0) == []
        # more keypoints than the context holds
        big = np.zeros((ctx.kp_cap + 1, 32), np.uint8)
        with pytest.raises(rgbdfe.RgbdfeError):
            ctx.projection_match(big, np.ones((len(big), 3), np.float32), None, f2["desc"], f2["xyz"], f2["pts_un"], proj, 5.0)
    finally:
        ctx.close()


@pytest.mark.gpu
def test_cuda_projection_match_device_frames(seq_vga):
    """Frames extracted on the device (distorted TUM fr1 camera) matched pairwise where they lie, against the oracle
    fed with the downloaded arrays; the pose of each pair comes from the device's own RansacSE3 result."""
    import frame_prep as fp
    import projection as pj
    import rgbdfe

    cam, frames = seq_vga
    cam = dict(cam, dist=[0.2624, -0.9531, -0.0054, 0.0026, 1.1633])
    n = len(frames)
    ctx = rgbdfe.Context(0, W, H, 1000, n)
    try:
        ctx.orb_configure(1000, 1.2, 8, 20, 7)
        fr = [rgbdfe.Frame(ctx) for _ in range(n)]
        bgr = [np.ascontiguousarray(np.stack([f[0]] * 3, 2)) for f in frames]
        d16 = [np.ascontiguousarray(np.clip(np.rint(f[1] * 5000.0), 0, 65535).astype(np.uint16)) for f in frames]
        ctx.frames_from_images(fr, bgr, d16, dict(cam, factor=5000.0))
        ctx.rng_seed(7)
        ctx.track_sequence(fr, 0)
        res = ctx.track_results(0, n - 1)
        bounds, inv = rgbdfe.camera_bounds(cam, W, H)
        assert bounds == tuple(np.float32(v) for v in fp.image_bounds(W, H, cam))
        host = []
        for f in fr:
            k, de, p3, outl = f.download()
            un, cell = f.download_undistorted()
            host.append(dict(desc=de, xyz=p3, outlier=outl.astype(bool), pts_un=un))
        projs, pairs = [], []
        eye = np.eye(3, dtype=np.float32)
        for i in range(n - 1):
            T21 = res[i]["T21"].reshape(4, 4).astype(np.float32)
            for th in (5.0, 15.0):
                projs.append(dict(fx=cam["fx"], fy=cam["fy"], cx=cam["cx"], cy=cam["cy"], Rwc1=eye, Ow1=np.zeros(3, np.float32),
                                  Rcw2=np.ascontiguousarray(T21[:3, :3]), tcw2=np.ascontiguousarray(T21[:3, 3]), bounds=bounds, inv=inv, th=th))
                pairs.append((i, i + 1))
        got = ctx.frames_projection_match([fr[a] for a, _ in pairs], [fr[b] for _, b in pairs], projs)
        total = 0
        for (a, b), p, g in zip(pairs, projs, got):
            ref = pj.projection_match(host[a], host[b], p, p["th"])
            assert _as_tuples(g) == ref, (a, b, p["th"])
            total += len(ref)
        assert total > 200, "the tracked pose must bring most keypoints onto their match"
    finally:
        ctx.close()
