#!/usr/bin/env python
"""bench.py — tracked frames/s of the RGB-D front end (ORB + match + SE3 RANSAC, 640x480).

Workload (BASELINE.json configs[1]): ICL-NUIM-shaped synthetic 640x480 sequence, 2000 ORB
features, kNN k=2 + ratio 0.9, RansacSE3 with 500 iterations.  One "step" = one batch of
BATCH consecutive frames: batched extraction (K1-K6) + per-frame match and RANSAC (K7-K11)
chained through the previous frame's outlier flags and the shared rand() stream.

  value : frames/s with the frames already resident in HBM (device pointers handed to the C ABI)
  e2e   : frames/s through the same C-ABI calls with pinned HOST buffers (H2D + result D2H timed)
Multi-GPU (--gpus N under torchrun): single-frame tracking does not shard (every frame needs the
previous frame's state), so ranks run independent replicas ("replicas only", weak scaling) and
additionally time the sharded loop-pair verification, reported under "verify".

--impl reference: the CPU oracle port (oracle/, the reference's algorithm restated; the reference
itself cannot be compiled here) on all host cores, same config / metric / unit.

The host loop is a rolling pipeline over 32 independent sequences (see Seq / run_group): results of a sequence's
previous batch are read just before its next batch is enqueued; the timed region ends with one drain.
Diagnostics (never set by the driver): --mode, --profile-host, --upload-mix, --small-ctas, --first-chunk, and the
environment knobs BENCH_HOSTPROF (host ms per sequence-step), BENCH_E2E_NO_DL / BENCH_E2E_DEPTH_DEV /
BENCH_E2E_GRAY_DEV (take one transport out of the e2e run), BENCH_VERIFY_CAP (hypothesis CTA shape of the verify leg).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
from concurrent.futures import ThreadPoolExecutor
import time

# Every sequence in flight owns two streams; with the default of 8 hardware work queues they alias and the chains
# of different sequences serialise behind each other's extraction kernels.  Must be set before CUDA initialises.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np

W, H, NFEAT, ITERS, MIN_INL, RATIO = 640, 480, 2000, 500, 10, 0.9
DEPTH_FACTOR = 5000.0  # DepthMapFactor of the TUM-format ICL-NUIM sequences
LEVEL_PX = [640 * 480, 533 * 400, 444 * 333, 370 * 278, 309 * 231, 257 * 193, 214 * 161, 179 * 134]
SUM_P = sum(LEVEL_PX)
# algorithmic bytes per frame of each extraction kernel (SURVEY 8d / DESIGN.md)
ALG_BYTES = {
    "k_pyramid_level(x levels)": LEVEL_PX[0] + sum(LEVEL_PX[1:]) + sum(LEVEL_PX[:-1]),
    "k_fast_cells": SUM_P,
    "k_gauss7": 2 * SUM_P,
    # gather set of K4+K6: every keypoint reads a 31 x 31 patch of its pyramid level (angle) and of the blurred level (descriptor);
    # 2000 overlapping patches cover the levels, so the set is bounded by the padded pyramid + the blurred levels of the frame
    "k_orient_describe": sum((w + 38 + 63) // 64 * 64 * (h + 38) + (w + 63) // 64 * 64 * h for w, h in
                             ((640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134))) + 60 * NFEAT,
    "k_knn2_umma": (NFEAT + NFEAT) * 32 + NFEAT * 16,  # both descriptor sets once + one int4 per query
}
# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of ONE launch over a 64-frame batch, from the ncu --set full
# captures summarised in profiles/r2_ncu_full_brief.txt, and the unit that actually limits each kernel there.
NCU_TRAFFIC_64 = {
    "k_fast_cells": 65.00e6 + 3.70e6, "k_gauss7": 72.88e6 + 32.24e6, "k_orient_describe": 139.05e6 + 9.58e6,
    "k_knn2_umma": 4.21e6, "k_quadtree": 8.17e6 + 0.01e6,
}
# warp instructions of one 64-frame launch (smsp__inst_executed.sum) -> thread instructions per pixel of the levels it covers
NCU_WARP_INST_64 = {"k_fast_cells": 173.76e6, "k_gauss7": 84.20e6, "k_orient_describe": 115.37e6, "k_quadtree": 31.92e6, "k_knn2_umma": 50.57e6}
NCU_LIMITER = {
    "k_fast_cells": "instruction-bound: issue slots 83 % busy, integer ALU pipe 74 % (16-bit-lane min/max; every integer ALU instruction costs a warp two issue clocks), tiles staged by TMA; DRAM traffic = 1.1x algorithmic, 5 % of peak",
    "k_gauss7": "instruction-bound: issue slots 86 % busy (IDP4A/IDP2A filter taps), L1/shared 60 %; DRAM 18 % of peak",
    "k_orient_describe": "XU 50 % (atan2 / sin / cos), L1 74 %, issue 73 %; DRAM 16 % of peak: the gather set of a 64-frame batch (141 MB) exceeds the 126 MB L2",
    "k_knn2_umma": "tcgen05.mma kind::i8 (UTCIMMA), tensor pipe 32 % busy; the warps that expand the operands and drain TMEM into the top-2 keep the integer ALU pipe 45 % busy and wait on each other at CTA barriers (2 CTAs x 8 warps per SM: TMEM-bound occupancy); DRAM < 1 % of peak",
    "k_quadtree": "latency (shared-memory atomics and scans, one CTA per level and frame): issue 32 %",
    "k_se3_hypotheses": "latency: one pair = 4-32 CTAs of dependent FP32 / FP64 chains; sm throughput 2 %, FP64 pipe 5 %, issue 14 % (ncu, profiles/r2_ncu_full_brief.txt)",
}
METRIC = "tracked frames/s (ORB+match+SE3 RANSAC, 640x480)"
WORKLOAD = "ICL-NUIM-shaped 640x480, 2000 ORB feats, kNN k=2 ratio 0.9, RansacSE3 500 iters (BASELINE configs[1])"


def make_sequence(n_frames):
    import synth

    cam = synth.CAMS["icl"]
    room = synth.Room(2345)
    poses = synth.xyz_trajectory(1000, 2347)[:n_frames]
    frames = []
    for R, t in poses:
        gray, depth = room.render(cam, R, t, 0.0, 0.0)
        # the dataset stores depth as 16-bit PNG (metres * DepthMapFactor); Frame converts it back with
        # convertTo(CV_32F, 1.0f / factor) (Core/Frame.cpp:48) -- both arms see exactly these values
        d16 = np.rint(np.clip(np.nan_to_num(depth, nan=0.0) * DEPTH_FACTOR, 0, 65535)).astype(np.uint16)
        frames.append((gray, d16.astype(np.float32) * (np.float32(1.0) / np.float32(DEPTH_FACTOR)), d16))
    return cam, frames


def pingpong(n):
    order = list(range(n)) + list(range(n - 2, 0, -1))
    i = 0
    while True:
        yield order[i % len(order)]
        i += 1


def sample_clocks_start():
    f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    try:
        p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200"], stdout=f, stderr=subprocess.DEVNULL)
    except Exception:
        return None, f.name
    return p, f.name


def sample_clocks_stop(p, path, gpu_index):
    if p is not None:
        time.sleep(0.25)
        p.terminate()
        try:
            p.wait(timeout=5)
        except Exception:
            p.kill()
    sm, mx, reasons = [], 0, set()
    try:
        for line in open(path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9 or c[0] != str(gpu_index):
                continue
            sm.append(float(c[1]))
            mx = max(mx, float(c[2]))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(path)
    except Exception:
        pass
    return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons)}


STAGES = ["pyramid", "fast", "quadtree", "angle", "blur", "describe", "knn", "filter", "ransac"]
OCV_STAGES = ("pyramid", "fast", "blur", "knn")  # the stages the reference runs inside OpenCV (resize, FAST, GaussianBlur, BFMatcher)


def native_oracle():
    """The oracle port rebuilt with the reference's own flags (-O3 -march=native, /root/reference CMakeLists.txt:6) ON THIS
    HOST (native code must not travel), keyed by the CPU model.  Timing only; parity uses the -O2 -ffp-contract=off build."""
    import hashlib

    try:
        model = [l for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        model = "unknown"
    rel = "_build/liboracle_native_%s.so" % hashlib.md5(model.encode()).hexdigest()[:8]
    path = os.path.join(ROOT, "oracle", rel)
    if not os.path.exists(path):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "native", "NATIVE_OUT=" + rel], stdout=sys.stderr)
    return C.CDLL(path), "-O3 -march=native"


def cpu_track(frames, cam, idx, threads=1, lib=None, want=False, cfg=None):
    """Time the oracle port over len(idx) frames per thread (one tracking chain per thread).
    Returns (tracked frames/s, seconds, mean per-thread stage seconds dict, (inliers, iters) of thread 0 if want)."""
    import orclib

    L = lib or orclib.lib()
    w_, h_, nfeat_, iters_, min_inl_ = cfg or (W, H, NFEAT, ITERS, MIN_INL)
    n = len(idx)
    g = (C.c_void_p * n)(*[frames[i][0].ctypes.data for i in idx])
    d = (C.c_void_p * n)(*[frames[i][1].ctypes.data for i in idx])
    done = [0] * threads
    stage = [(C.c_double * 16)() for _ in range(threads)]
    inl, its = (C.c_int * n)(), (C.c_int * n)()
    stats = (C.c_longlong * 6)()

    def work(t):
        done[t] = L.orc_track_sequence2(
            n, g, d, w_, h_, C.c_float(cam["fx"]), C.c_float(cam["fy"]), C.c_float(cam["cx"]), C.c_float(cam["cy"]), nfeat_, C.c_float(RATIO), iters_,
            C.c_uint(min_inl_), C.c_float(3.0), C.c_uint(4), C.c_uint(42 + t), inl if (want and t == 0) else None, its if (want and t == 0) else None, stage[t], stats if (want and t == 0) else None,
        )

    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    dt = time.perf_counter() - t0
    st = {name: float(np.mean([stage[t][i] for t in range(threads)])) for i, name in enumerate(STAGES)}
    res = (list(inl)[: n - 1], list(its)[: n - 1], list(stats)) if want else None
    return sum(done) / dt, dt, st, res


def cv2_stage_seconds(frames, n_frames, threads):
    """Seconds per frame (per thread, `threads` running at once) that the REAL OpenCV of this image (cv2, 1 thread per call)
    spends in the stages the reference hands to OpenCV: the 8-level resize + copyMakeBorder pyramid (ORBextractor.cpp:773-797),
    cv::FAST(20, nonmax) over every level (one call per level: a lower bound of the reference's 815 per-cell calls, :655-661),
    GaussianBlur 7x7 sigma 2 per level (:745-746) and BFMatcher(NORM_HAMMING).knnMatch(k=2) of 2000 x 2000 descriptors
    (Matcher.cpp:113).  Returns None when cv2 is missing."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    sizes = [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
    rng = np.random.default_rng(3)
    dq, dt_ = rng.integers(0, 256, (NFEAT, 32), dtype=np.uint8), rng.integers(0, 256, (NFEAT, 32), dtype=np.uint8)
    acc = [dict.fromkeys(OCV_STAGES, 0.0) for _ in range(threads)]

    def work(t):
        fast = cv2.FastFeatureDetector_create(20, True)
        bf = cv2.BFMatcher(cv2.NORM_HAMMING)
        a = acc[t]
        for k in range(n_frames):
            img = frames[(t + k) % len(frames)][0]
            t0 = time.perf_counter()
            lv = [cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)]
            for (w_, h_) in sizes[1:]:
                r = cv2.resize(lv[-1][19:-19, 19:-19], (w_, h_), interpolation=cv2.INTER_LINEAR)
                lv.append(cv2.copyMakeBorder(r, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))
            t1 = time.perf_counter()
            for im in lv:
                fast.detect(im[3:-3, 3:-3], None)
            t2 = time.perf_counter()
            for im in lv:
                cv2.GaussianBlur(im[19:-19, 19:-19], (7, 7), 2, 2, cv2.BORDER_REFLECT_101)
            t3 = time.perf_counter()
            bf.knnMatch(dq, dt_, k=2)
            t4 = time.perf_counter()
            a["pyramid"] += t1 - t0
            a["fast"] += t2 - t1
            a["blur"] += t3 - t2
            a["knn"] += t4 - t3

    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    return {k: float(np.mean([a[k] for a in acc])) / n_frames for k in OCV_STAGES}


def cpu_arm(frames, cam, n_frames, threads, lib, flags):
    """One pass of the CPU reference arm: `threads` tracking chains of n_frames frames each with the native build of the
    port, plus the same pass with the OpenCV stages replaced by what cv2 itself takes for them (hybrid: the reference links
    OpenCV for exactly those stages).  Returns dict(port=fps, hybrid=fps or None, ...)."""
    idx = [i % len(frames) for i in range(n_frames)]
    fps, dt, st, _ = cpu_track(frames, cam, idx, threads, lib)
    out = {"port_fps": fps, "seconds": dt, "port_stage_s_per_frame": {k: v / n_frames for k, v in st.items()}, "flags": flags}
    cvs = cv2_stage_seconds(frames, max(4, n_frames // 4), threads)
    if cvs:
        # per-thread wall seconds of the chain with the four OpenCV stages swapped for cv2's own time
        hybrid_dt = dt - sum(st[k] for k in OCV_STAGES) + sum(cvs[k] for k in OCV_STAGES) * n_frames
        best_dt = dt - sum(st[k] for k in OCV_STAGES) + sum(min(cvs[k] * n_frames, st[k]) for k in OCV_STAGES)
        out["cv2_stage_s_per_frame"] = cvs
        out["cv2_fps"] = fps * dt / hybrid_dt   # all four OpenCV stages at cv2's speed
        out["hybrid_fps"] = fps * dt / best_dt  # each OpenCV stage at the faster of cv2 and the port
    return out


def projection_leg(rgbdfe, ctx, fr, cam, B, no_cpu):
    """Row N2: Matcher::projectionMatch for the B-1 consecutive pairs of a device-resident batch, poses from the
    device's own RansacSE3 (identity pose for the first frame of each pair).  Host clock around the blocking call:
    four launches for all pairs, one count read-back in the middle, the matches copied to the host."""
    ctx.rng_seed(42)
    ctx.track_sequence(fr, 0)
    res = ctx.track_results(0, B - 1)
    bounds, inv = rgbdfe.camera_bounds(cam, W, H)
    eye = np.eye(3, dtype=np.float32)
    projs = []
    for i in range(B - 1):
        T = res[i]["T21"].reshape(4, 4).astype(np.float32)
        projs.append(rgbdfe.Projection.make(dict(fx=cam["fx"], fy=cam["fy"], cx=cam["cx"], cy=cam["cy"], Rwc1=eye, Ow1=np.zeros(3, np.float32),
                                                 Rcw2=T[:3, :3], tcw2=T[:3, 3], bounds=bounds, inv=inv), 5.0))
    f1, f2 = fr[:-1], fr[1:]
    got = None
    for _ in range(3):
        got = ctx.frames_projection_match(f1, f2, projs)
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        ctx.frames_projection_match(f1, f2, projs)
    dt = (time.perf_counter() - t0) / reps
    nm = [len(g) for g in got]
    out = {"value": (B - 1) / dt, "unit": "pairs/s", "ms_per_call": dt * 1e3, "pairs_per_call": B - 1,
           "matches_per_pair": float(np.mean(nm)), "claim_rounds_max": int(ctx.projection_rounds(B - 1).max()),
           "what": "rgbdfe_frames_projection_match, th = 5, device-resident frames (TUM fr1 distortion), matches copied to the host"}
    # one pair through the host-array entry point (what the Matcher shim calls)
    k0, d0, p0, o0 = fr[0].download()
    k1, d1, p1, _ = fr[1].download()
    un1, _ = fr[1].download_undistorted()
    t0 = time.perf_counter()
    for _ in range(20):
        ctx.projection_match(d0, p0, o0, d1, p1, un1, projs[0])
    out["host_arrays_ms_per_pair"] = (time.perf_counter() - t0) / 20 * 1e3
    if not no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import projection as pj

        T = res[0]["T21"].reshape(4, 4).astype(np.float32)
        pd = dict(fx=cam["fx"], fy=cam["fy"], cx=cam["cx"], cy=cam["cy"], Rwc1=eye, Ow1=np.zeros(3, np.float32), Rcw2=T[:3, :3], tcw2=T[:3, 3],
                  bounds=bounds, inv=inv)
        t0 = time.perf_counter()
        ref = pj.projection_match(dict(desc=d0, xyz=p0, outlier=o0.astype(bool)), dict(desc=d1, xyz=p1, pts_un=un1), pd, 5.0)
        out["cpu_oracle_pairs_per_s"] = 1.0 / (time.perf_counter() - t0)
        out["cpu_note"] = "numpy restatement (oracle/projection.py) on one host core, one pair; the reference never calls this function"
        out["parity_first_pair"] = [(int(a["queryIdx"]), int(a["trainIdx"]), float(a["distance"])) for a in got[0]] == ref
    return out


def cvorb_leg(rgbdfe, torch, local_rank, gray_d, depth_d, cam, inv_factor, B, nfeat, no_cpu, frames):
    """Row N3: Extractor(ORB, ORB) = cv::ORB with Harris scoring (Features/Extractor.cpp:55-59,163-166), batched
    extraction of device-resident frames through rgbdfe_frames_extract_raw after rgbdfe_cvorb_configure."""
    L = rgbdfe.lib()
    nseq = 4
    ctxs = [rgbdfe.Context(local_rank, W, H, nfeat, B) for _ in range(nseq)]
    frs = []
    for cx in ctxs:
        cx.cvorb_configure(nfeat, 1.2, 8)
        frs.append([rgbdfe.Frame(cx) for _ in range(B)])

    def step():
        for k, (cx, fr) in enumerate(zip(ctxs, frs)):
            idx = [(k * 17 + i) % gray_d.shape[0] for i in range(B)]
            fa = (C.c_void_p * B)(*[f.h for f in fr])
            gp = (C.c_void_p * B)(*[gray_d[i].data_ptr() for i in idx])
            dp = (C.c_void_p * B)(*[depth_d[i].data_ptr() for i in idx])
            cx._chk(L.rgbdfe_frames_extract_raw(cx.h, fa, B, gp, W, H, W, dp, W, C.c_float(inv_factor), C.c_float(cam["fx"]),
                                                C.c_float(cam["fy"]), C.c_float(cam["cx"]), C.c_float(cam["cy"])))

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 6
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    for cx in ctxs:
        cx.sync()
    dt = time.perf_counter() - t0
    out = {"value": steps * nseq * B / dt, "unit": "frames/s",
           "what": f"rgbdfe_cvorb_configure({nfeat}, 1.2, 8) + rgbdfe_frames_extract_raw: cv::ORB detect + compute + back-projection of "
                   f"device-resident 640x480 frames, {nseq} contexts x {B}-frame batches, host clock around enqueue + sync",
           "keypoints_per_frame": float(np.mean([len(frs[0][i].download()[0]) for i in range(4)]))}
    # per-kernel times of one batch (CUDA events on the context's stream)
    cx = ctxs[0]
    L.rgbdfe_profile_name.restype = C.c_char_p
    cx._chk(L.rgbdfe_profile_enable(cx.h, 1))
    for _ in range(2):
        step()
    a_ms, a_cnt = (C.c_double * 16)(), (C.c_longlong * 16)()
    cx._chk(L.rgbdfe_profile_read(cx.h, a_ms, a_cnt, 16))
    cx._chk(L.rgbdfe_profile_enable(cx.h, 0))
    names = {0: "k_cv_pyr_base+level(x7)", 1: "k_cv_fast", 2: "k_cv_blur", 3: "k_cv_select+gather", 4: "k_cv_orient_describe"}
    out["ms_per_batch"] = {names[i]: a_ms[i] / max(a_cnt[i], 1) for i in names}
    if not no_cpu:
        try:
            import cv2

            cv2.setNumThreads(1)
            det, des = cv2.ORB_create(nfeat, 1.2, 8), cv2.ORB_create()
            t0 = time.perf_counter()
            for i in range(20):
                k = det.detect(frames[i][0], None)
                des.compute(frames[i][0], k)
            out["cpu_opencv_frames_per_s"] = 20 / (time.perf_counter() - t0)
            out["cpu_note"] = "the real OpenCV (cv2 4.13, 1 thread): ORB_create(n, 1.2, 8).detect + ORB_create().compute, 20 frames"
        except Exception:
            pass
    for cx, fr in zip(ctxs, frs):
        for f in fr:
            f.close()
        cx.close()
    return out


def shim_leg(frames, cam, nfeat, n_frames=300):
    """The drop-in caller's path: ONE camera, blocking calls through the reference's class interfaces (Frame ctor ->
    Extractor::detectAndCompute, Matcher::match, RansacSE3::compute on host vectors, Tracking.cpp:121-163) in the C++
    caller tests/cpp/host_track.cpp, timed per class call with the host clock."""
    exe = os.path.join(ROOT, "rgbd-slam_b200", "lib", "host_track")
    with tempfile.TemporaryDirectory() as td:
        inp, outp = os.path.join(td, "in.bin"), os.path.join(td, "out.json")
        sub = frames[:24]
        with open(inp, "wb") as f:
            np.array([len(sub), W, H, nfeat], np.int32).tofile(f)
            np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32).tofile(f)
            for fr in sub:
                np.ascontiguousarray(fr[0], np.uint8).tofile(f)
                np.ascontiguousarray(fr[1], np.float32).tofile(f)
        subprocess.check_call([exe, inp, outp, "42", "orb2", "bench", str(n_frames)], stdout=sys.stderr)
        r = json.load(open(outp))
    r["what"] = "tests/cpp/host_track.cpp bench mode: Frame(gray, depth, extractor) + Matcher(0.9).match + RansacSE3(200,10,3,4).compute per frame, blocking, host vectors in and out"
    return r


def verify_leg(args, rgbdfe, torch, dist, rank, local_rank, world):
    """BASELINE configs[3] and [4] through the C ABI (rgbdfe_verify_sharded_enqueue / _collect, NCCL made by
    rgbdfe_nccl_comm_create): a FIXED job -- 1 query vs `--verify-kf` (4096) keyframes of 1000 features, the keyframes
    block-sharded over the ranks (strong scaling) -- and the batched relocalisation 256 queries x 1024 keyframes.
    Device time (CUDA events on the context stream around enqueue .. last collect), max over ranks.  Outside the timed
    region rank 0 recomputes pairs of every other rank's block on its own GPU and compares the gathered bytes."""
    import loopverify
    import synth

    N_FEAT = 1000
    cov = (0.01 * 2.0 * 2.0) ** 2

    def bcast_id():
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = torch.frombuffer(bytearray(rgbdfe.nccl_unique_id()), dtype=torch.uint8).cuda()
        if world > 1:
            dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    def timed_queries(vs, queries, n_queries_per_call, calls, depth=3, local_only=False):
        """`calls` enqueues of n_queries_per_call queries each, `depth` in flight, round-robin over the verifiers `vs` (contexts of this
        GPU: the latency-bound RANSAC tail of one call overlaps the kNN of the next); returns ms (device, max over ranks)."""
        vs = vs if isinstance(vs, (list, tuple)) else [vs]
        nv = len(vs)
        stream = torch.cuda.ExternalStream(vs[0].ctx.stream, device=local_rank)

        res_buf = np.zeros((n_queries_per_call, vs[0].n_kf), rgbdfe.SE3_DTYPE)  # every collect unpacks the full list into the same host array

        def run(n):
            for i in range(n + depth):
                if i >= depth:
                    j = i - depth
                    vs[j % nv].collect((j // nv) % 4, n_queries_per_call, out=res_buf)
                if i < n:
                    qd, qp = queries[i % len(queries)]
                    vs[i % nv].enqueue(qd, qp, N_FEAT, n_queries_per_call, (i // nv) % 4, seed=42)

        run(2 * nv)
        if world > 1 and not local_only:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = sum(v.ctx.launches for v in vs)
        e0.record(stream)
        run(calls)
        for v in vs[1:]:  # every slot was collected (all kernels, gathers and copies to the host have finished); order the event behind all streams
            vs[0].ctx.stream_wait(v.ctx)
        e1.record(stream)
        e1.synchronize()
        torch.cuda.synchronize()
        ms_ = e0.elapsed_time(e1)
        return (ms_ if local_only else max_over_ranks(ms_)), sum(v.ctx.launches for v in vs) - l0

    out = {"unit": "pairs/s", "features": N_FEAT, "api": "rgbdfe_verify_sharded_enqueue/_collect (C ABI), NCCL all-gather of 84 B/pair on the context's communication stream"}
    ident = bcast_id() if world > 1 else None

    # ---- configs[3]: 1 query vs n_kf keyframes, strong scaling ----
    n_kf = args.verify_kf
    qd, qp, kd, kp, _ = synth.synth_pair_sets(n_kf, N_FEAT, seed=5678, overlap_frac=0.05)  # the same data on every rank
    lo, hi = rgbdfe.shard_range(n_kf, rank, world)
    v = loopverify.CudaShardedVerifier(local_rank, rank, world, n_kf, kd[lo:hi], kp[lo:hi], ident, max_feats=N_FEAT, depth_cov=cov)
    queries = [(np.ascontiguousarray(np.roll(qd, 13 * i, axis=0)), np.ascontiguousarray(np.roll(qp, 13 * i, axis=0))) for i in range(8)]
    vcap = os.environ.get("BENCH_VERIFY_CAP")  # diagnostic: hypothesis CTA shape of the many-pairs path (option ransac_verify_cap)
    if vcap is not None:
        v.ctx.set_option("ransac_verify_cap", int(vcap))
    # `value`: the query frames are resident in HBM (device pointers are read in place); the host-memory form is timed beside it
    q_dev = [(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()) for a, b in queries]
    q_ptr = [(a.data_ptr(), b.data_ptr()) for a, b in q_dev]
    calls = 48
    ms_host, _ = timed_queries(v, queries, 1, calls)
    ms_one, launches = timed_queries(v, q_ptr, 1, calls)
    # two contexts of this GPU take alternate queries (each its own stream, workspace, resident keyframe block and communicator)
    vb = loopverify.CudaShardedVerifier(local_rank, rank, world, n_kf, kd[lo:hi], kp[lo:hi], bcast_id() if world > 1 else None, max_feats=N_FEAT, depth_cov=cov)
    if vcap is not None:
        vb.ctx.set_option("ransac_verify_cap", int(vcap))
    ms, _ = timed_queries([v, vb], q_ptr, 1, calls, depth=4)
    vb.close()
    out["host_queries_pairs_per_s"] = calls * n_kf / (ms_host * 1e-3)
    out["one_context_pairs_per_s"] = calls * n_kf / (ms_one * 1e-3)
    out["contexts_per_gpu"] = 2
    out.update({"value": calls * n_kf / (ms * 1e-3), "keyframes_total": n_kf, "keyframes_per_gpu": hi - lo, "ms_per_query": ms / calls,
                "scaling": "strong (fixed 1 query x %d keyframes job, keyframes block-sharded over %d GPU(s))" % (n_kf, world),
                "launches_per_query": launches / calls})
    # parity on hardware: gathered results vs rank 0's own GPU recomputing runs of every block (all pairs at N <= 2)
    full = v.verify(queries[0][0], queries[0][1], seed=42)[0]
    parity, checked = True, 0
    if rank == 0:
        run = n_kf if world <= 2 else max(64, 512 // world)
        for r in range(world):
            rlo, rhi = rgbdfe.shard_range(n_kf, r, world)
            rhi = min(rhi, rlo + run)
            db = v.ctx.kfdb_create(kd[rlo:rhi], kp[rlo:rhi])
            ref = v.ctx.verify_batch(db, rhi - rlo, queries[0][0], queries[0][1], 0.9, 30, 200, 30, 3.0, 4, seed=42, first_pair_index=rlo)
            v.ctx.kfdb_destroy(db)
            parity = parity and ref.tobytes() == full[rlo:rhi].tobytes()
            checked += rhi - rlo
    out.update({"parity": bool(parity), "parity_pairs_checked": checked, "loops_found": int(full["ok"].sum()),
                "parity_how": "rank 0 recomputed runs of every rank's block alone with rgbdfe_verify_batch; gathered bytes must be identical"})
    v.close()
    if world > 1:
        dist.barrier()
    # the same fixed job on ONE GPU (rank 0 alone), for the strong-scaling efficiency of this very run
    if world > 1:
        single = None
        if rank == 0:
            v1 = [loopverify.CudaShardedVerifier(local_rank, 0, 1, n_kf, kd, kp, None, max_feats=N_FEAT, depth_cov=cov) for _ in range(2)]
            try:
                ms1, _ = timed_queries(v1, q_ptr, 1, 16, depth=4, local_only=True)
                ms1 /= 16
            finally:
                for x_ in v1:
                    x_.close()
            single = n_kf / (ms1 * 1e-3)
        dist.barrier()
        if rank == 0:
            out["one_gpu_same_job_pairs_per_s"] = single
            out["strong_efficiency"] = out["value"] / (world * single)
    else:
        out["strong_efficiency"] = 1.0

    # ---- configs[4]: batched relocalisation, 256 queries x 1024 keyframes, one gather per call ----
    try:
        Q, n_kf4, per_call = 256, 1024, 32
        qd4, qp4, kd4, kp4, _ = synth.synth_pair_sets(n_kf4, N_FEAT, seed=777, overlap_frac=0.05)
        lo, hi = rgbdfe.shard_range(n_kf4, rank, world)
        v4 = [loopverify.CudaShardedVerifier(local_rank, rank, world, n_kf4, kd4[lo:hi], kp4[lo:hi], bcast_id() if world > 1 else None,
                                             max_feats=N_FEAT, depth_cov=cov) for _ in range(2)]
        qs = np.stack([np.roll(qd4, 7 * i, axis=0) for i in range(per_call)])
        ps = np.stack([np.roll(qp4, 7 * i, axis=0) for i in range(per_call)])
        reps4 = 3  # three 256-query batches back to back: start-up and tail of the two-context pipeline amortised
        qs_d, ps_d = torch.from_numpy(qs).cuda(), torch.from_numpy(ps).cuda()  # query frames resident in HBM, as for configs[3]
        ms4, _ = timed_queries(v4, [(qs_d.data_ptr(), ps_d.data_ptr())], per_call, reps4 * Q // per_call, depth=4)
        out["relocalisation"] = {"value": reps4 * Q * n_kf4 / (ms4 * 1e-3), "unit": "pairs/s", "queries": Q, "keyframes_total": n_kf4,
                                 "keyframes_per_gpu": hi - lo, "ms_per_256_query_batch": ms4 / reps4,
                                 "what": "BASELINE configs[4]: 256 query frames x 1024 keyframes, all pairs match + RansacSE3(200, 30), "
                                         "%d queries per call, one all-gather per call, two contexts per GPU taking alternate calls" % per_call}
        for x_ in v4:
            x_.close()
    except Exception as ex:
        out["relocalisation"] = {"error": str(ex)[:200]}
    return out


def config_leg(rgbdfe, torch, local_rank, cam_name, nfeat, iters, min_inl, n_frames, B_, S_, steps, no_cpu, what, first_chunk=4):
    """Another BASELINE config through the same pipeline as the headline (batched extraction of device-resident frames in
    one context, the match / RANSAC chain in another, S_ independent sequences in flight): frames/s on the device (CUDA
    events), mean executed RANSAC iterations, and the CPU port (reference flags, all cores) on the same frames beside it."""
    import synth

    cam, fr_ = synth.sequence(cam_name, n=n_frames, seed=4321, noise_sigma=0.0, invalid_frac=0.0)
    w_, h_ = cam["w"], cam["h"]
    grays = [np.ascontiguousarray(g) for g, _ in fr_]
    d16 = [np.rint(np.clip(np.nan_to_num(d, nan=0.0) * DEPTH_FACTOR, 0, 65535)).astype(np.uint16) for _, d in fr_]
    inv = float(np.float32(1.0) / np.float32(DEPTH_FACTOR))
    gd, dd = torch.from_numpy(np.stack(grays)).cuda(), torch.from_numpy(np.stack(d16)).cuda()
    L = rgbdfe.lib()

    class Q:
        def __init__(self, k):
            self.cx, self.ct = rgbdfe.Context(local_rank, w_, h_, nfeat, B_), rgbdfe.Context(local_rank, w_, h_, nfeat, 1)
            self.cx.orb_configure(nfeat, 1.2, 8, 20, 7)
            self.ct.set_stream_priority(True)
            self.ct.set_option("ransac_first_chunk", first_chunk if S_ > 1 else 32)
            self.ct.rng_seed(42 + k)
            self.sets = [[rgbdfe.Frame(self.cx) for _ in range(B_)] for _ in range(3)]
            self.order, self.i, self.have_ref = pingpong(n_frames), 0, False
            for _ in range((k * 5) % n_frames):
                next(self.order)
            self.ok = self.n = 0
            self.iters, self.inl = [], []

        def extract(self, batch):
            idx = [next(self.order) for _ in range(B_)]
            fa = (C.c_void_p * B_)(*[f.h for f in batch])
            gp = (C.c_void_p * B_)(*[gd[i].data_ptr() for i in idx])
            dp = (C.c_void_p * B_)(*[dd[i].data_ptr() for i in idx])
            self.cx._chk(L.rgbdfe_frames_extract_raw(self.cx.h, fa, B_, gp, w_, h_, w_, dp, w_, C.c_float(inv), C.c_float(cam["fx"]), C.c_float(cam["fy"]),
                                                     C.c_float(cam["cx"]), C.c_float(cam["cy"])))

        def step(self):
            i = self.i
            cur, nxt = self.sets[i % 3], self.sets[(i + 1) % 3]
            self.ct.stream_wait(self.cx)
            self.extract(nxt)
            chain = ([self.sets[(i - 1) % 3][-1]] if self.have_ref else []) + cur
            self.ct.track_sequence(chain, 0, RATIO, True, iters, min_inl, 3.0, 4, True)
            self.n_res = len(chain) - 1

        def collect(self):
            res = self.ct.track_results(0, self.n_res)
            self.ok += int(res["ok"].sum())
            self.n += len(res)
            self.iters.append(float(res["iters_executed"].mean()))
            self.inl.append(float(res["n_inliers"].mean()))
            self.i += 1
            self.have_ref = True

        def close(self):
            for st in self.sets:
                for f in st:
                    f.close()
            self.cx.close()
            self.ct.close()

    qs = [Q(k) for k in range(S_)]
    try:
        for q in qs:
            q.extract(q.sets[0])

        def run(n):
            for _ in range(n):
                for q in qs:
                    q.step()
                for q in qs:
                    q.collect()

        run(2)
        torch.cuda.synchronize()
        stream = torch.cuda.ExternalStream(qs[0].ct.stream, device=local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for q in qs:
            q.iters, q.inl, q.ok, q.n = [], [], 0, 0
        e0.record(stream)
        run(steps)
        for q in qs:
            qs[0].ct.stream_wait(q.cx)
            qs[0].ct.stream_wait(q.ct)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        out = {"what": what, "value": steps * S_ * B_ / (ms * 1e-3), "unit": "frames/s", "image": f"{w_}x{h_}", "features": nfeat, "ransac": f"RansacSE3({iters}, {min_inl}, 3.0, 4)",
               "sequences_in_flight": S_, "batch": B_, "ransac_first_chunk": first_chunk, "mean_ransac_iterations": float(np.mean([x for q in qs for x in q.iters])),
               "mean_inliers": float(np.mean([x for q in qs for x in q.inl])), "tracked_ok_frac": sum(q.ok for q in qs) / max(sum(q.n for q in qs), 1),
               "keypoints_per_frame": int(len(qs[0].sets[0][0].download()[0]))}
    finally:
        for q in qs:
            q.close()
    if not no_cpu:
        cores = len(os.sched_getaffinity(0))
        nlib, flags = native_oracle()
        fl = [(grays[i], d16[i].astype(np.float32) * (np.float32(1.0) / np.float32(DEPTH_FACTOR))) for i in range(n_frames)]
        nf = min(n_frames, 9)
        fps, dt, st, _ = cpu_track(fl, cam, list(range(nf)), cores, nlib, cfg=(w_, h_, nfeat, iters, min_inl))
        out["cpu_port_frames_per_s"] = fps
        out["cpu_note"] = f"oracle port built {flags}, {nf}-frame chains on {cores} threads, {dt:.1f} s"
    return out


def hard_ransac_leg(rgbdfe, torch, local_rank, no_cpu):
    """RansacSE3 where it has to work: 256 pairs of 2000-feature frames whose descriptor matches are right but 70 % of the
    matched 3-D points are wrong (moved at random), RansacSE3(500, 10, 3.0, 4): the loop runs hundreds of iterations per pair
    instead of breaking after the first.  Pairs/s through rgbdfe_verify_sharded (one GPU), executed iterations from the results,
    and the oracle port on a few of the same pairs beside it."""
    import loopverify
    import synth

    n_pairs, nf = 256, 2000
    qd, qp, kd, kp, _ = synth.synth_pair_sets(n_pairs, nf, seed=99, corr_frac=0.5, overlap_frac=1.0)
    rng = np.random.default_rng(5)
    bad = rng.random((n_pairs, nf)) < 0.7
    z = rng.uniform(0.5, 4.5, (n_pairs, nf))
    junk = np.stack([rng.uniform(-1, 1, (n_pairs, nf)) * z, rng.uniform(-0.8, 0.8, (n_pairs, nf)) * z, z], 2).astype(np.float32)
    kp = np.where(bad[:, :, None], junk, kp)
    v = loopverify.CudaShardedVerifier(local_rank, 0, 1, n_pairs, kd, kp, None, max_feats=nf)
    v.params = rgbdfe.VerifyParams(0.9, 10, 500, 10, 3.0, 4, 42)
    try:
        res = v.verify(qd, qp, seed=42)[0]
        stream = torch.cuda.ExternalStream(v.ctx.stream, device=local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 6
        torch.cuda.synchronize()
        e0.record(stream)
        for i in range(reps):
            v.enqueue(qd, qp, nf, 1, i % 4, seed=42)
            if i >= 3:
                v.collect((i - 3) % 4, 1, want=False)
        for i in range(reps - 3, reps):
            v.collect(i % 4, 1, want=False)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1) / reps
    finally:
        v.close()
    out = {"what": "hard RANSAC: 256 pairs x 2000 features, ~1000 matches per pair of which ~70 % are geometric outliers, RansacSE3(500, 10, 3.0, 4)",
           "value": n_pairs / (ms * 1e-3), "unit": "pairs/s", "ms_per_256_pairs": ms, "mean_ransac_iterations": float(res["iters_executed"].mean()),
           "mean_matches": float(res["n_matches"].mean()), "mean_inliers": float(res["n_inliers"].mean()), "ok_frac": float((res["ok"] != 0).mean())}
    if not no_cpu:
        import orclib

        t0 = time.perf_counter()
        its = []
        for k in range(4):
            m = orclib.match(kd[k], qd, np.zeros(nf, np.uint8), kp[k], qp, 0.9, True)
            r_, l_ = orclib.Rng(42 + k), orclib.Latch()
            l_.latched.value, l_.cov.value = 1, (0.01 * 2.0 * 2.0) ** 2
            r = orclib.ransac_se3(kp[k], qp, m, r_, l_, 500, 10, 3.0, 4)
            its.append(r["iters_executed"])
            out["parity_first_pairs"] = bool(out.get("parity_first_pairs", True) and r["iters_executed"] == int(res[k]["iters_executed"])
                                             and len(r["inliers"]) == int(res[k]["n_inliers"]))
        out["cpu_port_pairs_per_s_one_thread"] = 4 / (time.perf_counter() - t0)
        out["cpu_note"] = "oracle parity build (-O2), match + RansacSE3 of the first 4 pairs, one thread"
    return out


def run_reference(args, rank, world):
    """The CPU arm: the reference's algorithm on all host cores.  The reference itself cannot be built here (OpenCV / Eigen
    / PCL headers are absent), so this is (a) the oracle port compiled with the reference's flags and (b) the same with the
    OpenCV stages at the speed of the real OpenCV of this image; `value` is the FASTER of the two."""
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    per_thread = 41  # one 41-frame chain per thread and step: 40 tracked frames, the first extraction amortised as on the GPU arm
    cam, frames = make_sequence(per_thread)
    lib, flags = native_oracle()
    for _ in range(min(args.warmup, 1)):
        cpu_arm(frames, cam, 5, cores, lib, flags)
    t0 = time.perf_counter()
    port, hybrid, cvh, last = [], [], [], None
    for _ in range(args.steps):
        last = cpu_arm(frames, cam, per_thread, cores, lib, flags)
        port.append(last["port_fps"])
        hybrid.append(last.get("hybrid_fps") or 0.0)
        cvh.append(last.get("cv2_fps") or 0.0)
    total = time.perf_counter() - t0
    v_port, v_hyb, v_cv = float(np.mean(port)), float(np.mean(hybrid)), float(np.mean(cvh))
    v = max(v_port, v_hyb, v_cv)
    kind = "port"
    sample = (f"{per_thread}-frame chains x {cores} threads per step; value = the fastest of: oracle port built {flags} ({v_port:.1f} frames/s), the same "
              f"with pyramid/FAST/blur/kNN at the speed of cv2's own resize/FAST/GaussianBlur/knnMatch ({v_cv:.1f}), and with each of those four "
              f"stages at the faster of cv2 and the port ({v_hyb:.1f})")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32/f64",
        "data": "synthetic", "config": {"workload": WORKLOAD, "path": "cpu (oracle port; the reference does not build in this image)",
                                        "port_frames_per_s": v_port, "port_with_cv2_stages_frames_per_s": v_cv, "port_with_faster_of_both_per_stage_frames_per_s": v_hyb,
                                        "port_stage_ms_per_frame": {k: 1e3 * x for k, x in last["port_stage_s_per_frame"].items()},
                                        "cv2_stage_ms_per_frame": {k: 1e3 * x for k, x in (last.get("cv2_stage_s_per_frame") or {}).items()}},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--frames", type=int, default=96)
    ap.add_argument("--sequences", type=int, default=32, help="independent tracked sequences in flight on each GPU")
    ap.add_argument("--extract-streams", type=int, default=0, help="extraction contexts share this many streams (0 = one stream each)")
    ap.add_argument("--overlap-blur", type=int, default=0, help="diagnostic: run the blur beside the quadtree on a side stream")
    ap.add_argument("--knn-waves", type=int, default=0, help="diagnostic: waves of kNN CTAs the train-range split aims for (0 = library default)")
    ap.add_argument("--first-chunk", type=int, default=1, help="speculative RANSAC iterations of the first launch when several sequences share the GPU "
                    "(1: no speculation; measured 72.9 k / 69.6 k / 64.8 k / 56.7 k frames/s at 1 / 2 / 4 / 8: idle speculative CTAs take SM space from the throughput kernels)")
    ap.add_argument("--host-threads", type=int, default=2, help="host threads that enqueue the sequences' work")
    ap.add_argument("--full-upload", action="store_true", help="upload the whole depth map of every frame instead of reading the keypoint sectors in place")
    ap.add_argument("--upload-mix", type=int, default=3, help="e2e: of every 8 batches of a sequence, this many upload the whole pinned depth map (copy engine); "
                    "the others read the keypoint sectors in place (PCIe reads): the two transports have different limits")
    ap.add_argument("--small-ctas", type=int, default=1, help="register-capped 128 / 256-thread hypothesis CTAs for the tracking chains (they leave room on their SM for the throughput kernels; 0 = 512-thread CTAs)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--profile-host", action="store_true", help="diagnostic: the per-kernel timing pass runs with pinned HOST inputs (the e2e path)")
    ap.add_argument("--core-only", action="store_true", help="profiling runs: only the headline workload, none of the other legs")
    ap.add_argument("--latency-sms", type=int, default=0,
                    help="SMs given to the match/RANSAC chains (green-context partition); 0 = share all SMs through stream priorities")
    ap.add_argument("--mode", default="full", choices=["full", "extract", "chain", "serial", "verify", "shim", "configs"],
                    help="diagnostic: time only the extraction half or only the match/RANSAC chains (not a bench value)")
    ap.add_argument("--verify-kf", type=int, default=4096, help="keyframes of the fixed loop-closure verification job (BASELINE configs[3]), sharded over the GPUs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import rgbdfe

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL prints its version banner on stdout when NCCL_DEBUG is set in the environment: stdout carries the one JSON
        # line only, so file descriptor 1 points at stderr while the communicator comes up (first collective included)
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            warm = torch.ones(1, device="cuda")
            dist.all_reduce(warm)
            g0 = torch.empty(world, device="cuda")
            dist.all_gather_into_tensor(g0, warm / world)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    if args.mode == "verify":  # diagnostic: the sharded verification leg alone
        out = verify_leg(args, rgbdfe, torch, dist, rank, local_rank, world)
        if rank == 0:
            print(json.dumps({"n_gpus": world, "verify": out}))
        if world > 1:
            dist.destroy_process_group()
        return

    B, F = args.batch, args.frames
    cam, frames = make_sequence(F)
    if args.mode == "configs":  # diagnostic: the other-config legs alone
        o = {"configs0": config_leg(rgbdfe, torch, local_rank, "tum_fr1", 1000, 200, 10, 32, 32, 32, 4, args.no_cpu, "configs[0]"),
             "configs2": config_leg(rgbdfe, torch, local_rank, "kinect2", 4000, 500, 10, 8, 8, 8, 4, args.no_cpu, "configs[2]", first_chunk=32),
             "hard": hard_ransac_leg(rgbdfe, torch, local_rank, args.no_cpu)}
        print(json.dumps(o))
        return
    if args.mode == "shim":  # diagnostic: the class-shim leg alone
        print(json.dumps({"class_shim_path": {"feats_2000": shim_leg(frames, cam, NFEAT), "feats_1000": shim_leg(frames, cam, 1000)}}))
        return
    gray_h = torch.from_numpy(np.stack([f[0] for f in frames])).pin_memory()
    depth_h = torch.from_numpy(np.stack([f[2] for f in frames])).pin_memory()  # raw 16-bit depth, as read from the PNG
    inv_factor = float(np.float32(1.0) / np.float32(DEPTH_FACTOR))
    gray_d, depth_d = gray_h.cuda(), depth_h.cuda()
    gp_h, dp_h = gray_h.numpy(), depth_h.numpy()
    L = rgbdfe.lib()
    # addresses of the pool's frames, taken once: building them per call costs the enqueuing thread 0.1 ms per batch
    gray_d_ptr, depth_d_ptr = [gray_d[i].data_ptr() for i in range(F)], [depth_d[i].data_ptr() for i in range(F)]
    gray_h_ptr, depth_h_ptr = [gp_h[i].ctypes.data for i in range(F)], [dp_h[i].ctypes.data for i in range(F)]
    c_cam = [C.c_float(inv_factor)] + [C.c_float(cam[k]) for k in ("fx", "fy", "cx", "cy")]

    part = rgbdfe.device_partition(local_rank, args.latency_sms) if args.latency_sms > 0 else None

    stream_owners = []

    class Seq:
        """One tracked sequence: two contexts of one device — ctx_x extracts frame batch n+1 (throughput-bound, big
        grids) while ctx_t runs the sequential match/RANSAC chain of batch n on a high-priority stream;
        rgbdfe_stream_wait orders them.  Its own rand() stream and depth-covariance latch, like one reference process."""

        def __init__(self, k, first_chunk):
            self.ctx_x = rgbdfe.Context(local_rank, W, H, NFEAT, B)
            self.ctx_t = rgbdfe.Context(local_rank, W, H, NFEAT, 1)
            self.ctx_x.orb_configure(NFEAT, 1.2, 8, 20, 7)
            self.full_upload = args.full_upload
            if self.full_upload:
                self.ctx_x.set_option("host_zero_copy", 0)
            self.ctx_x.set_option("depth_upload_of_8", args.upload_mix)
            self.ctx_x.set_option("depth_upload_phase", k % 8)  # the sequences' upload turns are staggered
            if args.overlap_blur:
                self.ctx_x.set_option("overlap_blur", 1)
            if args.extract_streams and k >= args.extract_streams:
                self.ctx_x.share_stream(stream_owners[k % args.extract_streams])
            else:
                stream_owners.append(self.ctx_x)
            if part:
                self.ctx_x.bind_partition(rgbdfe.PART_THROUGHPUT)
                self.ctx_t.bind_partition(rgbdfe.PART_LATENCY)
            else:
                self.ctx_t.set_stream_priority(True)
            self.ctx_t.set_option("ransac_first_chunk", first_chunk)
            if args.small_ctas:
                self.ctx_t.set_option("ransac_small_ctas", 1)
            if args.knn_waves:
                self.ctx_t.set_option("knn_waves", args.knn_waves)
            self.ctx_t.rng_seed(42 + k)
            self.sets = [[rgbdfe.Frame(self.ctx_x) for _ in range(B)] for _ in range(3)]
            self.fr_arr = [(C.c_void_p * B)(*[f.h for f in st]) for st in self.sets]
            self.chain_ref = [(C.c_void_p * (B + 1))(*([self.sets[(k - 1) % 3][-1].h] + [f.h for f in self.sets[k]])) for k in range(3)]
            self.order = pingpong(F)
            for _ in range((k * 37) % F):  # sequences start at different frames
                next(self.order)
            self.i, self.have_ref, self.n_res, self.pending = 0, False, 0, False
            self.inliers, self.ok, self.n, self.iters = [], 0, 0, []
            # e2e: what the class interface hands back per frame -- keypoints + descriptors (Extractor::detectAndCompute) and
            # mvInliers (RansacSE3) -- lands in pinned host buffers, one set per frame set
            cap = self.ctx_x.kp_cap
            self.h_kps = [torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory().numpy() for _ in range(3)]
            self.h_desc = [torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory().numpy() for _ in range(3)]
            self.h_n = [torch.zeros(B, dtype=torch.int32).pin_memory().numpy() for _ in range(3)]
            self.d2h_inlier_records = 0
            self.pending_download = False
            self.r_res = np.zeros(B + 1, rgbdfe.SE3_DTYPE)
            self.r_inl = np.zeros((B + 1) * self.ctx_t.kp_cap, rgbdfe.DM_DTYPE)
            self.r_off = np.zeros(B + 2, np.int32)

        def extract(self, k, dev):
            """Extraction of the next 64 frames of the sequence into frame set k: device-resident inputs read in place, or
            pinned host inputs (gray uploaded, depth read over PCIe) -- the same C-ABI call either way."""
            idx = [next(self.order) for _ in range(B)]
            cx = self.ctx_x
            gp = (C.c_void_p * B)(*[(gray_d_ptr if (dev or os.environ.get("BENCH_E2E_GRAY_DEV")) else gray_h_ptr)[i] for i in idx])  # (diagnostic knob)
            dp = (C.c_void_p * B)(*[(depth_d_ptr if (dev or os.environ.get("BENCH_E2E_DEPTH_DEV")) else depth_h_ptr)[i] for i in idx])  # (diagnostic knob)
            cx._chk(L.rgbdfe_frames_extract_raw(cx.h, self.fr_arr[k], B, gp, W, H, W, dp, W, *c_cam))

        def enqueue(self, dev):
            i = self.i
            self.ctx_t.stream_wait(self.ctx_x)  # batch i was extracted by the previous step
            if args.mode != "chain":
                k = (i + 1) % 3
                self.extract(k, dev)            # batch i+1, concurrent with the chain below
                if not dev and not os.environ.get("BENCH_E2E_NO_DL"):  # e2e: the frames' host-side state comes back too (asynchronous, pinned)
                    self.ctx_x._chk(L.rgbdfe_frames_download_async(self.ctx_x.h, self.fr_arr[k], B, rgbdfe._p(self.h_kps[k]), rgbdfe._p(self.h_desc[k]),
                                                                   None, self.ctx_x.kp_cap, rgbdfe._p(self.h_n[k])))
                    self.pending_download = True
            if args.mode == "serial":           # diagnostic: the chain waits for this extraction too (no overlap)
                self.ctx_t.stream_wait(self.ctx_x)
            if args.mode == "extract":
                self.n_res = 0
                return
            chain, n = (self.chain_ref[i % 3], B + 1) if self.have_ref else (self.fr_arr[i % 3], B)
            self.ctx_t._chk(L.rgbdfe_track_sequence(self.ctx_t.h, chain, n, C.c_float(RATIO), 1, ITERS, C.c_uint(MIN_INL), C.c_float(3.0), C.c_uint(4), 1, 0))
            self.n_res = n - 1

        def collect(self):
            if self.n_res == 0:
                self.i += 1
                return
            if self.pending_download:  # e2e: results WITH the inlier sets, and the downloaded frames complete
                res = self.r_res[: self.n_res]
                self.ctx_t._chk(L.rgbdfe_track_results_inliers(self.ctx_t.h, 0, self.n_res, rgbdfe._p(res), rgbdfe._p(self.r_inl), len(self.r_inl),
                                                               rgbdfe._p(self.r_off)))
                self.d2h_inlier_records += int(self.r_off[self.n_res])
                self.ctx_x.sync()
                self.pending_download = False
            else:
                res = self.ctx_t.track_results(0, self.n_res)
            self.ok += int(res["ok"].sum())
            self.n += len(res)
            self.inliers.append(float(res["n_inliers"].mean()))
            self.iters.append(float(res["iters_executed"].mean()))
            self.i += 1
            self.have_ref = True

        def close(self):
            for st in self.sets:
                for f in st:
                    f.close()
            self.ctx_x.close()
            self.ctx_t.close()

    pool = ThreadPoolExecutor(max_workers=args.host_threads) if args.host_threads > 1 else None

    host_prof, host_prof_on = {}, [False]

    def run_group(group, dev, steps):
        """A rolling pipeline: a sequence's results of step n are read just before its step n+1 is enqueued, so the other
        sequences' work keeps the GPU busy while the host waits for one chain; everything is drained at the end."""
        prof = host_prof.setdefault((dev, len(group)), [0.0, 0.0, 0]) if (os.environ.get("BENCH_HOSTPROF") and host_prof_on[0]) else None
        for _ in range(steps):
            for q in group:
                t0 = time.perf_counter()
                if q.pending:
                    q.collect()
                t1 = time.perf_counter()
                q.enqueue(dev)
                q.pending = True
                if prof is not None:
                    prof[0] += t1 - t0
                    prof[1] += time.perf_counter() - t1
                    prof[2] += 1
        for q in group:
            q.collect()
            q.pending = False

    def run_steps(seqs, dev, steps):
        """Every sequence advances by `steps` batches (extraction of the next + chain of the current, then its results).
        The C-ABI calls release the GIL, so a few host threads can feed the 2 x len(seqs) streams concurrently."""
        T = args.host_threads
        if pool and len(seqs) > 1:
            list(pool.map(lambda k: run_group(seqs[k::T], dev, steps), range(T)))
        else:
            run_group(seqs, dev, steps)
        return B * len(seqs) * steps

    def timed(seqs, dev, steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stream = torch.cuda.ExternalStream(seqs[0].ctx_t.stream, device=local_rank)
        l0 = sum(q.ctx_x.launches + q.ctx_t.launches for q in seqs)
        e0.record(stream)
        host_prof_on[0] = True
        n = run_steps(seqs, dev, steps)
        host_prof_on[0] = False
        for q in seqs:  # the last step's extraction (and every other sequence) belongs to the timed region
            seqs[0].ctx_t.stream_wait(q.ctx_x)
            seqs[0].ctx_t.stream_wait(q.ctx_t)
        e1.record(stream)
        e1.synchronize()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return n, ms, sum(q.ctx_x.launches + q.ctx_t.launches for q in seqs) - l0

    S = max(1, args.sequences)
    seqs = [Seq(k, 32 if S == 1 else args.first_chunk) for k in range(S)]
    for q in seqs:
        q.extract(0, True)  # prime the pipelines
    for _ in range(args.warmup):
        run_steps(seqs, True, 1)
        run_steps(seqs, False, 1)
    proc, path = sample_clocks_start() if rank == 0 else (None, None)
    n_dev, ms_dev, launches = timed(seqs, True, args.steps)
    for q in seqs:
        q.d2h_inlier_records = 0
    n_e2e, ms_e2e, _ = timed(seqs, False, args.steps)
    e2e_inlier_records = sum(q.d2h_inlier_records for q in seqs)
    # the same with the whole 16-bit depth map uploaded per frame (option host_zero_copy = 0)
    for q in seqs:
        q.ctx_x.set_option("host_zero_copy", 0)
    run_steps(seqs, False, 1)
    n_up, ms_up, _ = timed(seqs, False, max(2, args.steps // 2))
    for q in seqs:
        q.ctx_x.set_option("host_zero_copy", 0 if q.full_upload else 1)
    clocks = sample_clocks_stop(proc, path, local_rank) if rank == 0 else None
    for dev_, (tc, te, n_) in host_prof.items():  # BENCH_HOSTPROF=1: host seconds per sequence-step in collect (incl. waiting) / enqueue
        print(f"hostprof dev={dev_} collect {tc / n_ * 1e3:.3f} ms enqueue {te / n_ * 1e3:.3f} ms per sequence-step ({n_})", file=sys.stderr)
    value = n_dev * world / (ms_dev * 1e-3)
    up_frac = 1.0 if args.full_upload else args.upload_mix / 8.0
    e2e = n_e2e * world / (ms_e2e * 1e-3)
    # one sequence alone (latency-bound reference point)
    single = None
    if S > 1:
        seqs[0].ctx_t.set_option("ransac_first_chunk", 32)
        seqs[0].ctx_t.set_option("ransac_small_ctas", 0)  # one camera alone: 512-thread hypothesis CTAs, 32 speculative iterations
        run_steps(seqs[:1], True, 1)
        n1, ms1, _ = timed(seqs[:1], True, max(2, args.steps // 2))
        single = n1 / (ms1 * 1e-3)
        seqs[0].ctx_t.set_option("ransac_first_chunk", args.first_chunk)
        seqs[0].ctx_t.set_option("ransac_small_ctas", args.small_ctas)
    ctx_x, ctx_t = seqs[0].ctx_x, seqs[0].ctx_t
    state = {"ok": sum(q.ok for q in seqs), "n": sum(q.n for q in seqs), "inliers": [x for q in seqs for x in q.inliers],
             "iters": [x for q in seqs for x in q.iters]}

    def step(dev):
        return run_steps(seqs[:1], dev, 1)

    # ---- per-kernel timing pass (CUDA events around every kernel, separate from the headline timing) ----
    lib = rgbdfe.lib()
    lib.rgbdfe_profile_name.restype = C.c_char_p
    ms_k = [0.0] * 16
    cnt_k = [0] * 16
    for cx in (ctx_x, ctx_t):
        cx._chk(lib.rgbdfe_profile_enable(cx.h, 1))
    prof_steps = 2
    for _ in range(prof_steps):
        step(not args.profile_host)
    for cx in (ctx_x, ctx_t):
        a_ms = (C.c_double * 16)()
        a_cnt = (C.c_longlong * 16)()
        cx._chk(lib.rgbdfe_profile_read(cx.h, a_ms, a_cnt, 16))
        cx._chk(lib.rgbdfe_profile_enable(cx.h, 0))
        for k in range(16):
            ms_k[k] += a_ms[k]
            cnt_k[k] += a_cnt[k]
    kernels = {lib.rgbdfe_profile_name(i).decode(): (ms_k[i], cnt_k[i]) for i in range(10)}
    tot_ms = sum(v[0] for v in kernels.values()) or 1.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    # dominant HBM-bound kernel of the extraction stage
    # (the pyramid is 8 launches of two kernels per batch and is listed under per_kernel only)
    ext = {k: v for k, v in kernels.items() if k in ALG_BYTES and v[1] > 0 and not k.startswith("k_pyramid")}
    if not ext:  # --mode chain (diagnostic): no extraction kernel ran
        ext = {k: v for k, v in kernels.items() if v[1] > 0}
    dom = max(ext, key=lambda k: ext[k][0])
    dom_ms, dom_cnt = ext[dom]
    per_launch_bytes = ALG_BYTES.get(dom, 0) * B
    achieved = per_launch_bytes / (dom_ms / dom_cnt * 1e-3) / 1e9
    traffic = NCU_TRAFFIC_64.get(dom)
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
        "traffic": (traffic * B / 64.0) if traffic else None,
        "algorithmic_bytes_per_launch": per_launch_bytes,
        "limiter": NCU_LIMITER.get(dom),
        "thread_instructions_per_pixel": (NCU_WARP_INST_64[dom] * 32 / (64.0 * SUM_P)) if dom in NCU_WARP_INST_64 else None,
        "bound_note": "the kernel is instruction-bound, not memory-bound: `frac` says how far it is from the HBM roof, `limiter` and the instructions per pixel say what to cut",
        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
        "per_kernel": {k: {"ms": v[0] / v[1], "algorithmic_GBps": ALG_BYTES[k] * B / (v[0] / v[1] * 1e-3) / 1e9} for k, v in kernels.items()
                       if k in ALG_BYTES and v[1] > 0},
        "share_of_one_sequence_kernel_time": {k: round(v[0] / tot_ms, 4) for k, v in kernels.items()},
        "share_note": "summed CUDA-event time of ONE sequence's kernels run alone (its 64-pair chain is serial, so the two k_se3 kernels dominate); in the step the chains of 32 sequences overlap the throughput kernels",
        "ms_per_launch": {k: (v[0] / v[1] if v[1] else None) for k, v in kernels.items()},
    }

    # the one tensor-pipe kernel: u8 multiply-accumulates of the 2000 x 256 by 256 x 2000 product of every pair against the
    # dense int8 rate (2 x the measured bf16 rate of MEASURED_PEAKS.json; nominal 4500 TOP/s if that file is absent)
    kn = roofline["per_kernel"].get("k_knn2_umma")
    if kn:
        tops = 2.0 * NFEAT * NFEAT * 256 * B / (kn["ms"] * 1e-3) / 1e12  # a batch's chain = B pairs of NFEAT x NFEAT descriptors
        peak_tops, peak_how = 2.0 * float(peaks.get("bf16_tflops", 2250.0)), "2 x measured dense bf16 (MEASURED_PEAKS.json)"
        try:  # the measured dispatch rate of tcgen05.mma.kind::i8 on this pool's B200 (tools/micro/umma_rate.cu)
            um = json.load(open(os.path.join(ROOT, "profiles", "r2_pipe_peaks.json")))["umma_i8_m128n128k32"]
            peak_tops = 2.0 * um["mac_per_clk_per_sm"] * 148 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12
            peak_how = "measured tcgen05.mma.kind::i8 dispatch rate (8181 MAC/clk/SM, profiles/r2_pipe_peaks.json) x 148 SMs x max SM clock"
        except Exception:
            pass
        kn.update({"bound": "tensor", "achieved_TOPs": tops, "peak_TOPs": peak_tops, "frac": tops / peak_tops,
                   "how": "2 * Nq * Nt * 256 * pairs per launch / launch time; peak = " + peak_how})

    # ---- row N1: Frame preprocessing (BGR -> gray, 16-bit depth, undistortion, grid cells) + extraction, end to end ----
    frame_prep = projection = None
    try:
        if args.core_only:
            raise RuntimeError("skipped (--core-only)")
        tum = dict(fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989, dist=[0.262383, -0.953104, -0.005358, 0.002628, 1.163314],
                   factor=DEPTH_FACTOR)
        nseq = 8
        bgr_h = torch.from_numpy(np.stack([np.stack([np.roll(f[0], k, 1) for k in (0, 3, 7)], 2) for f in frames[:B]])).pin_memory()
        pctx = [rgbdfe.Context(local_rank, W, H, NFEAT, B) for _ in range(nseq)]
        pfr = []
        for cx_ in pctx:
            cx_.orb_configure(NFEAT, 1.2, 8, 20, 7)
            pfr.append([rgbdfe.Frame(cx_) for _ in range(B)])
        bn, dn = bgr_h.numpy(), depth_h.numpy()

        def pstep():
            for cx_, fr_ in zip(pctx, pfr):
                cx_.frames_from_images(fr_, [bn[i] for i in range(B)], [dn[i] for i in range(B)], tum)

        for _ in range(2):
            pstep()
        for cx_ in pctx:
            cx_.sync()
        t0 = time.perf_counter()
        psteps = 4
        for _ in range(psteps):
            pstep()
        for cx_ in pctx:
            cx_.sync()
        dt = time.perf_counter() - t0
        frame_prep = {"value": psteps * nseq * B / dt, "unit": "frames/s",
                      "what": "rgbdfe_frames_from_images: pinned BGR u8 + 16-bit depth -> gray, pyramid, ORB, undistorted keypoints, grid cells, 3D points "
                              "(TUM fr1 distortion), 8 contexts x 64-frame batches, host clock around enqueue + sync",
                      "h2d_bytes_per_frame": W * H * 3 + NFEAT * 32}
        try:
            import cv2

            K = np.array([[tum["fx"], 0, tum["cx"]], [0, tum["fy"], tum["cy"]], [0, 0, 1]], np.float32)
            D = np.array(tum["dist"], np.float32)
            pts = np.random.default_rng(0).uniform(0, 480, (NFEAT, 1, 2)).astype(np.float32)
            t0 = time.perf_counter()
            for i in range(20):
                cv2.cvtColor(bn[i], cv2.COLOR_BGR2GRAY)
                dn[i].astype(np.float32) * np.float32(1.0 / DEPTH_FACTOR)
                cv2.undistortPoints(pts, K, D, None, K)
            frame_prep["cpu_preprocessing_only_frames_per_s"] = 20 / (time.perf_counter() - t0)
            frame_prep["cpu_note"] = "cv2 cvtColor + depth scale + undistortPoints alone on one host core (no extraction)"
        except Exception:
            pass
        # ---- row N2: Matcher::projectionMatch on the frames one of those contexts just produced ----
        try:
            projection = projection_leg(rgbdfe, pctx[0], pfr[0], tum, B, args.no_cpu)
        except Exception as ex:
            projection = {"error": str(ex)[:200]}
        for cx_, fr_ in zip(pctx, pfr):
            for f in fr_:
                f.close()
            cx_.close()
    except Exception as ex:
        frame_prep = {"error": str(ex)[:200]}

    # ---- row N3: the cv::ORB branch of Extractor ----
    try:
        if args.core_only:
            raise RuntimeError("skipped (--core-only)")
        cvorb = cvorb_leg(rgbdfe, torch, local_rank, gray_d, depth_d, cam, inv_factor, B, NFEAT, args.no_cpu, frames)
    except Exception as ex:
        cvorb = {"error": str(ex)[:200]}

    # ---- sharded loop-pair verification (the part of the path that shards), N >= 1 ----
    try:
        if args.core_only:
            raise RuntimeError("skipped (--core-only)")
        verify = verify_leg(args, rgbdfe, torch, dist, rank, local_rank, world)
    except Exception as ex:  # the headline metric does not depend on this leg
        verify = {"error": str(ex)[:300]}

    # ---- the other BASELINE configs and a hard-RANSAC workload, same pipeline (rank 0 only; not part of the headline) ----
    other = {}
    if rank == 0 and args.mode == "full" and not args.core_only:
        for key, a in (("configs0_tum_1000", ("tum_fr1", 1000, 200, 10, 32, 32, 32, 4, "BASELINE configs[0]: TUM fr1-shaped 640x480, 1000 ORB features, RansacSE3(200, 10) as Tracking.cpp:129")),
                       ("configs2_1080p_4000", ("kinect2", 4000, 500, 10, 8, 8, 8, 4, "BASELINE configs[2]: Kinect-v2-shaped 1920x1080 frames with registered depth, 4000 ORB features"))):
            try:
                other[key] = config_leg(rgbdfe, torch, local_rank, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], args.no_cpu, a[8],
                                        first_chunk=32 if key.startswith("configs2") else 4)  # configs[2]'s pairs run hundreds of iterations: 32 hypotheses at a time
            except Exception as ex:
                other[key] = {"error": str(ex)[:200]}
        try:
            other["hard_ransac"] = hard_ransac_leg(rgbdfe, torch, local_rank, args.no_cpu)
        except Exception as ex:
            other["hard_ransac"] = {"error": str(ex)[:200]}

    # ---- the class-shim path (what a drop-in user of the reference's classes gets): one camera, blocking calls ----
    shim = None
    if rank == 0 and not args.core_only:
        try:
            shim = {"feats_2000": shim_leg(frames, cam, NFEAT), "feats_1000": shim_leg(frames, cam, 1000)}
        except Exception as ex:
            shim = {"error": str(ex)[:200]}

    cpu = parity = None
    if rank == 0 and not args.no_cpu:
        cores = len(os.sched_getaffinity(0))
        nlib, flags = native_oracle()
        arm = cpu_arm(frames, cam, 21, cores, nlib, flags)
        best = max(arm["port_fps"], arm.get("hybrid_fps") or 0.0, arm.get("cv2_fps") or 0.0)
        cpu = {"value": best, "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": f"21-frame chains (20 tracked frames) on each of {cores} threads, {arm['seconds']:.1f} s: the fastest of the oracle port built {flags} "
                         f"({arm['port_fps']:.1f} frames/s), the same with pyramid/FAST/blur/kNN at the speed of cv2's own resize/FAST/GaussianBlur/knnMatch "
                         f"({(arm.get('cv2_fps') or 0.0):.1f}) and with each of those stages at the faster of cv2 and the port ({(arm.get('hybrid_fps') or 0.0):.1f})",
               "port_stage_ms_per_frame": {k: 1e3 * x for k, x in arm["port_stage_s_per_frame"].items()},
               "cv2_stage_ms_per_frame": {k: 1e3 * x for k, x in (arm.get("cv2_stage_s_per_frame") or {}).items()}}
        # ---- parity inside the bench run: a fresh sequence (srand(42), frames 0..40) on the GPU against the parity build of the oracle ----
        try:
            n_par = min(41, F)
            pcx, pct = rgbdfe.Context(local_rank, W, H, NFEAT, 64), rgbdfe.Context(local_rank, W, H, NFEAT, 1)
            pcx.orb_configure(NFEAT, 1.2, 8, 20, 7)
            pct.rng_seed(42)
            pfr_ = [rgbdfe.Frame(pcx) for _ in range(n_par)]
            pcx.frames_extract_raw(pfr_, [gp_h[i] for i in range(n_par)], [dp_h[i] for i in range(n_par)], cam, inv_factor)
            pcx.sync()
            pct.track_sequence(pfr_, 0, RATIO, True, ITERS, MIN_INL, 3.0, 4, True)
            gres = pct.track_results(0, n_par - 1)
            _, odt, _, (o_inl, o_its, o_stats) = cpu_track(frames, cam, list(range(n_par)), 1, None, want=True)
            same_inl = [int(x) for x in gres["n_inliers"]] == o_inl
            same_its = [int(x) for x in gres["iters_executed"]] == o_its
            parity = {"parity_checked": bool(same_inl and same_its), "pairs": n_par - 1, "inlier_counts_equal": bool(same_inl),
                      "iteration_counts_equal": bool(same_its), "how": "fresh sequence, srand(42), frames 0..%d: n_inliers and executed RANSAC iterations of every pair, GPU (C ABI) vs the oracle's parity build (-O2 -ffp-contract=off), single thread %.1f s" % (n_par - 1, odt),
                      "oracle_se3_ops_per_pair": {k: o_stats[i] / max(o_stats[5], 1) for i, k in enumerate(("E_s", "E_f", "A", "V", "rounds"))}}
            for f in pfr_:
                f.close()
            pcx.close()
            pct.close()
        except Exception as ex:
            parity = {"parity_checked": False, "error": str(ex)[:200]}
        # ---- RANSAC kernels against the measured FP32 / FP64 pipe rates (tools/micro/pipe_peaks.cu -> profiles/r2_pipe_peaks.json) ----
        try:
            pk = json.load(open(os.path.join(ROOT, "profiles", "r2_pipe_peaks.json")))
            ops = (parity or {}).get("oracle_se3_ops_per_pair")
            t_hyp, t_prep = roofline["ms_per_launch"].get("k_se3_hypotheses"), roofline["ms_per_launch"].get("k_se3_prepare")
            if ops and t_hyp:
                f64 = 29 * ops["E_s"] + 170 * ops["E_f"]   # flops of errorFunction2 (shortcut / full)
                f32 = 40 * ops["A"] + 400 * ops["V"]       # TFC add / 3x3 SVD + transform
                peak64 = 2 * pk["dfma_f64"]["device_Gops_per_s"]  # an FMA counts 2 flops
                peak32 = 2 * pk["ffma_f32"]["device_Gops_per_s"]
                roofline["ransac"] = {
                    "kernel": "k_se3_hypotheses (+ k_se3_prepare)", "bound": "latency (dependent FP32 / FP64 chains of one pair; one chain per sequence)",
                    "ms_per_pair": {"k_se3_prepare": t_prep, "k_se3_hypotheses": t_hyp}, "flops_per_pair": {"f64": f64, "f32": f32},
                    "achieved_GFLOPs": {"f64": f64 / (t_hyp * 1e-3) / 1e9, "f32": f32 / (t_hyp * 1e-3) / 1e9},
                    "peak_GFLOPs": {"f64_dfma": peak64, "f32_ffma": peak32, "source": "profiles/r2_pipe_peaks.json (tools/micro/pipe_peaks.cu on this pool's B200)"},
                    "frac": {"f64": f64 / (t_hyp * 1e-3) / 1e9 / peak64, "f32": f32 / (t_hyp * 1e-3) / 1e9 / peak32},
                    "note": "one pair alone on the device: a single chain cannot fill a pipe; the step hides the chains of the other sequences behind the throughput kernels",
                    "ncu": NCU_LIMITER.get("k_se3_hypotheses")}
        except Exception:
            pass

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev / args.steps, "frames_per_step": S * B, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32/f64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "sequences_in_flight": S, "single_sequence_frames_per_s": single, "mean_ransac_iterations": float(np.mean(state["iters"])),
                       "batch_frames_per_step": B, "distinct_frames": F, "pipeline": "extraction of batch n+1 (context A) overlaps the match/RANSAC chain of batch n (context B)",
                       "sm_partition": ({"latency_sms": part[0], "throughput_sms": part[1], "how": "CUDA green contexts"} if part else "none (stream priorities)"), "input": "gray u8 + raw 16-bit depth per frame (Frame's convertTo(CV_32F, 1/5000) runs on the device)",
                       "host_threads": args.host_threads,
                       "l2": f"inputs larger than L2: {S} sequences x rotating {F}-frame set ({F * (W * H * 3) / 1e6:.0f} MB) + 58 MB of pyramid per batch",
                       "multi_gpu": "replicas only (tracking does not shard); sharded verification under 'verify'",
                       "tracked_ok_frac": state["ok"] / max(state["n"], 1), "mean_inliers": float(np.mean(state["inliers"])),
                       "parity_checked": (parity or {}).get("parity_checked"),
                       "configs0_tum_1000_feats_frames_per_s": (other.get("configs0_tum_1000") or {}).get("value"),
                       "configs2_1080p_4000_feats_frames_per_s": (other.get("configs2_1080p_4000") or {}).get("value"),
                       "hard_ransac_pairs_per_s": (other.get("hard_ransac") or {}).get("value"),
                       "hard_ransac_mean_iterations": (other.get("hard_ransac") or {}).get("mean_ransac_iterations"),
                       "class_shim_frames_per_s_2000_feats": ((shim or {}).get("feats_2000") or {}).get("frames_per_s"),
                       "class_shim_frames_per_s_1000_feats": ((shim or {}).get("feats_1000") or {}).get("frames_per_s"),
                       # the sharded leg (BASELINE configs[3] / [4]); details under the top-level key "verify"
                       "verify_pairs_per_s": (verify or {}).get("value"), "verify_keyframes_total": (verify or {}).get("keyframes_total"),
                       "verify_ms_per_query": (verify or {}).get("ms_per_query"), "verify_strong_efficiency": (verify or {}).get("strong_efficiency"),
                       "verify_one_gpu_same_job_pairs_per_s": (verify or {}).get("one_gpu_same_job_pairs_per_s"),
                       "verify_parity": (verify or {}).get("parity"), "verify_parity_pairs_checked": (verify or {}).get("parity_pairs_checked"),
                       "reloc_pairs_per_s": ((verify or {}).get("relocalisation") or {}).get("value"), "verify_error": (verify or {}).get("error")},
            "e2e": {"value": e2e, "unit": "frames/s",
                    "h2d_bytes_per_step": int(S * B * (W * H * (1 + 2 * up_frac) + (1 - up_frac) * NFEAT * 32)),
                    "d2h_bytes_per_step": S * B * (seqs[0].ctx_x.kp_cap * 60 + 4 + 84 + 4) + e2e_inlier_records * 16 // max(args.steps, 1),
                    "d2h_note": "per frame: keypoints (28 B) + descriptors (32 B) for the frame's whole capacity row, the count, the 84-byte RansacSE3 result and mvInliers (16 B per inlier DMatch) -- what Extractor::detectAndCompute and RansacSE3 hand back through the class interface",
                    "h2d_note": ("gray image uploaded per frame; the pinned 16-bit depth map is uploaded whole for %d of every 8 batches (copy engine) and read in place for the others (only the 32 B sectors under the <= 2000 keypoints cross PCIe: k_unproject_host) -- the two transports have different limits" % args.upload_mix) if not args.full_upload else "gray + whole 16-bit depth map uploaded per frame",
                    "depth_transport": {"upload_batches_of_8": 8 if args.full_upload else args.upload_mix,
                                        "why": "with the depth map resident the same run reaches the device-resident value; in-place reads alone sustain ~146 M sector reads/s, whole uploads alone are PCIe-bandwidth-bound (full_upload_value)"},
                    "full_upload_value": n_up / (ms_up * 1e-3) * world, "full_upload_h2d_bytes_per_step": S * B * W * H * 3},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "class_shim_path": shim, "other_configs": other, "verify": verify, "frame_prep": frame_prep, "projection_match": projection, "cvorb_extract": cvorb,
        }))
    for q in reversed(seqs):  # contexts that share a stream close before its owner
        q.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
