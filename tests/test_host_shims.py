"""The reference's class interfaces (Extractor, ORBextractor, Matcher, RansacSE3, Ransac, Random) as host shims:
a C++ caller written like Tracking::visualOdometry (tests/cpp/host_track.cpp) is run on the GPU box and its
outputs are compared with the CPU oracle, bit for bit (poses within 1e-5)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "rgbd-slam_b200", "lib")


def fnv(a):
    h = 1469598103934665603
    for b in np.ascontiguousarray(a).view(np.uint8).ravel().tolist():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def test_host_library_exports_the_reference_classes():
    so = os.path.join(LIB, "librgbdfe_host.so")
    assert os.path.exists(so), "build it with __graft_entry__.build()"
    syms = subprocess.check_output(["nm", "-DC", "--defined-only", so], text=True)
    for want in (
        "Extractor::Extractor(Extractor::eType, Extractor::eType, Extractor::eMode)",
        "Extractor::setParameters(int, float, int, int, int)",
        "Extractor::detectAndCompute(",
        "Extractor::getScaleFactors()",
        "Extractor::mNorm",
        "ORBextractor::ORBextractor(int, float, int, int, int)",
        "ORBextractor::operator()(",
        "Matcher::Matcher(float)",
        "Matcher::match(std::shared_ptr<Frame>, std::shared_ptr<Frame>, std::vector<cv::DMatch",
        "Matcher::projectionMatch(std::shared_ptr<Frame>, std::shared_ptr<Frame>, std::vector<cv::DMatch",
        "Matcher::descriptorDistance(cv::Mat const&, cv::Mat const&)",
        "RansacSE3::RansacSE3(int, unsigned int, float, unsigned int)",
        "RansacSE3::compute(std::shared_ptr<Frame>, std::shared_ptr<Frame>, std::vector<cv::DMatch",
        "RansacSE3::setIterations(int)",
        "Ransac::Ransac(std::shared_ptr<Frame>, std::shared_ptr<Frame>, std::vector<cv::DMatch",
        "Ransac::compute(std::vector<cv::DMatch",
        "Random::randomInt(int, int)",
        "Random::initSeed()",
    ):
        assert want in syms, want


@pytest.mark.gpu
def test_visual_odometry_through_the_class_shims(oracle, seq_vga, tmp_path):
    import synth

    cam, frames = seq_vga
    inp, outp = tmp_path / "in.bin", tmp_path / "out.txt"
    with open(inp, "wb") as f:
        np.array([len(frames), 640, 480, 1000], np.int32).tofile(f)
        np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32).tofile(f)
        for g, d in frames:
            np.ascontiguousarray(g, np.uint8).tofile(f)
            np.ascontiguousarray(d, np.float32).tofile(f)
    exe = os.path.join(LIB, "host_track")
    subprocess.check_call([exe, str(inp), str(outp), "42"])
    lines = open(outp).read().splitlines()
    rec = {}
    for ln in lines:
        t = ln.split()
        rec.setdefault(t[0], []).append(t[1:])
    assert rec["levels"][0][0] == "8" and rec["levels"][0][4] == "6"  # NORM_HAMMING
    orb = oracle.Orb(1000)
    tab = orb.tables()
    assert np.float32(rec["scale7"][0][0]) == np.float32(tab["scale"][7]) and np.float32(rec["scale7"][0][2]) == np.float32(tab["inv_sigma2"][7])
    rng, latch = oracle.Rng(42), oracle.Latch()
    prev, pose_prev = None, np.eye(4, dtype=np.float32)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    import orclib
    import projection as pj

    orclib_dm, nproj = orclib.DM_DTYPE, 0
    ti = 0
    for i, (g, d) in enumerate(frames):
        k, de = orb.extract(g)
        p3 = synth.unproject(cam, np.stack([k["x"], k["y"]], 1), d)
        fr = rec["frame"][i]
        assert int(fr[2]) == len(k)
        assert int(fr[4], 16) == fnv(k) and int(fr[6], 16) == fnv(de) and int(fr[8], 16) == fnv(p3)
        outl = np.zeros(len(k), np.uint8)
        if prev is not None:
            m = oracle.match(prev[1], de, prev[3], prev[2], p3, 0.9, True)
            r = oracle.ransac_se3(prev[2], p3, m, rng, latch, 200, 10, 3.0, 4)
            if len(m) >= 10:
                outl[m["trainIdx"]] = 1
            if r["ok"]:
                outl[r["inliers"]["trainIdx"]] = 0
            # Matcher::descriptorDistance (row a13): Hamming distance of 32-byte rows, recomputed from the oracle's descriptors
            dd = rec["ddist"][ti]
            ham = lambda a, b: float(np.unpackbits(a ^ b).sum())
            assert int(dd[0]) == i
            assert [float(x) for x in dd[1:4]] == [ham(prev[1][0], de[0]), ham(prev[1][1], de[-1]), 0.0]
            tr = rec["track"][ti]
            assert int(tr[2]) == len(m) and int(tr[4], 16) == fnv(m)
            assert int(tr[6]) == int(r["ok"]) and int(tr[8]) == len(r["inliers"]) and int(tr[10], 16) == fnv(r["inliers"])
            assert np.float32(tr[12]) == r["rmse"]
            T = np.array([np.float32(x) for x in tr[14:30]], np.float32).reshape(4, 4)
            assert np.abs(T - r["T"]).max() <= 1e-5 and (T == r["T"]).all()
            pose = np.array([np.float32(x) for x in rec["pose"][ti][1:17]], np.float32).reshape(4, 4)
            want = (r["T"] @ pose_prev) if r["ok"] else np.eye(4, dtype=np.float32)
            assert np.abs(pose - want).max() <= 1e-5
            pose_prev = pose
            assert int(rec["outliers"][ti][1]) == int(outl.sum())
            # Matcher::projectionMatch through the shim against the numpy oracle, with the pose blocks the shim read
            pr = rec["proj"][ti]
            blk = np.array([np.float32(x) for x in pr[6:30]], np.float32)
            proj = dict(fx=cam["fx"], fy=cam["fy"], cx=cam["cx"], cy=cam["cy"], Rwc1=blk[0:9].reshape(3, 3), Ow1=blk[9:12], Rcw2=blk[12:21].reshape(3, 3),
                        tcw2=blk[21:24], bounds=(0.0, 640.0, 0.0, 480.0), inv=(np.float32(64) / np.float32(640), np.float32(48) / np.float32(480)))
            f1 = dict(desc=prev[1], xyz=prev[2], outlier=prev[3].astype(bool))
            f2 = dict(desc=de, xyz=p3, pts_un=np.stack([k["x"], k["y"]], 1).astype(np.float32))
            pm = pj.projection_match(f1, f2, proj, 5.0)
            dm = np.zeros(len(pm), orclib_dm)
            if pm:
                dm["queryIdx"], dm["trainIdx"], dm["distance"] = (np.array(v) for v in zip(*pm))
            dm["imgIdx"] = -1
            assert int(pr[2]) == len(pm) and int(pr[4], 16) == fnv(dm)
            nproj += len(pm)
            if i == 1:
                u = oracle.ransac_umeyama(prev[2], p3, m, rng, 0, 0.02, 0.2, 10, 3)
                ur = rec["umeyama"][0]
                assert int(ur[1]) == int(u["ok"]) and int(ur[3]) == len(u["inliers"]) and int(ur[5], 16) == fnv(u["inliers"])
                assert int(rec["randomInt"][0][0]) == rng.random_int(0, 999)
            ti += 1
        prev = (k, de, p3, outl)
    assert nproj > 100, "the tracked poses must let projectionMatch find most of the matches"


@pytest.mark.gpu
def test_extractor_orb_orb_pair_through_the_class_shims(oracle, seq_vga, tmp_path):
    """Extractor(ORB, ORB) (cv::ORB with Harris scoring, Features/Extractor.cpp:55-59,163-166,216-218) through the
    same caller: features, Matcher::match and RansacSE3 against the oracle fed with the oracle's cv::ORB features."""
    import synth

    cam, frames = seq_vga
    inp, outp = tmp_path / "in.bin", tmp_path / "out.txt"
    with open(inp, "wb") as f:
        np.array([len(frames), 640, 480, 1000], np.int32).tofile(f)
        np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32).tofile(f)
        for g, d in frames:
            np.ascontiguousarray(g, np.uint8).tofile(f)
            np.ascontiguousarray(d, np.float32).tofile(f)
    subprocess.check_call([os.path.join(LIB, "host_track"), str(inp), str(outp), "42", "orb"])
    rec = {}
    for ln in open(outp).read().splitlines():
        t = ln.split()
        rec.setdefault(t[0], []).append(t[1:])
    assert rec["levels"][0][0] == "8" and rec["levels"][0][4] == "6"  # nlevels, NORM_HAMMING
    assert np.float32(rec["levels"][0][2]) == np.float32(np.float64(np.float32(1.2)))  # getScaleFactor(): the double member
    s7 = np.float32(1.0)
    for _ in range(7):  # Extractor::init (Extractor.cpp:63-80): float = float * double
        s7 = np.float32(np.float64(s7) * np.float64(np.float32(1.2)))
    assert np.float32(rec["scale7"][0][0]) == s7
    orb = oracle.CvOrb(1000, 1.2, 8)
    rng, latch = oracle.Rng(42), oracle.Latch()
    prev, ti = None, 0
    for i, (g, d) in enumerate(frames):
        k, de = orb.extract(g)
        p3 = synth.unproject(cam, np.stack([k["x"], k["y"]], 1), d)
        fr = rec["frame"][i]
        assert int(fr[2]) == len(k)
        assert int(fr[4], 16) == fnv(k) and int(fr[6], 16) == fnv(de) and int(fr[8], 16) == fnv(p3)
        outl = np.zeros(len(k), np.uint8)
        if prev is not None:
            m = oracle.match(prev[1], de, prev[3], prev[2], p3, 0.9, True)
            r = oracle.ransac_se3(prev[2], p3, m, rng, latch, 200, 10, 3.0, 4)
            if len(m) >= 10:
                outl[m["trainIdx"]] = 1
            if r["ok"]:
                outl[r["inliers"]["trainIdx"]] = 0
            tr = rec["track"][ti]
            assert int(tr[2]) == len(m) and int(tr[4], 16) == fnv(m)
            assert int(tr[6]) == int(r["ok"]) and int(tr[8]) == len(r["inliers"]) and int(tr[10], 16) == fnv(r["inliers"])
            T = np.array([np.float32(x) for x in tr[14:30]], np.float32).reshape(4, 4)
            assert np.abs(T - r["T"]).max() <= 1e-5
            if i == 1:  # the caller also runs class Ransac and Random here: keep the oracle's rand() stream in step
                oracle.ransac_umeyama(prev[2], p3, m, rng, 0, 0.02, 0.2, 10, 3)
                rng.random_int(0, 999)
            ti += 1
        prev = (k, de, p3, outl)


@pytest.mark.gpu
def test_shim_path_timing_mode_runs(seq_vga, tmp_path):
    """host_track's bench mode (the drop-in caller's blocking loop, timed per class call) produces sane numbers."""
    import json

    cam, frames = seq_vga
    inp, outp = tmp_path / "in.bin", tmp_path / "out.json"
    with open(inp, "wb") as f:
        np.array([len(frames), 640, 480, 1000], np.int32).tofile(f)
        np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32).tofile(f)
        for g, d in frames:
            np.ascontiguousarray(g, np.uint8).tofile(f)
            np.ascontiguousarray(d, np.float32).tofile(f)
    subprocess.check_call([os.path.join(LIB, "host_track"), str(inp), str(outp), "42", "orb2", "bench", "40"])
    r = json.load(open(outp))
    assert r["frames"] == 40 and r["frames_per_s"] > 50 and r["ok_frac"] > 0.9 and r["mean_inliers"] > 100
