#!/usr/bin/env python
"""Generates the golden vectors under tests/golden/ (run once, in the build container; outputs are committed).

Three files:

  frame_prep_cv2.npz   cv2 4.13 outputs for Frame's preprocessing (row N1): cvtColor(BGR2GRAY) of a colour image and
                       cv::undistortPoints of 3000 points with the TUM fr1 intrinsics / distortion.

  cv2_primitives.npz   outputs of the REAL third-party code the reference calls on this path, produced here with
                       cv2 4.13 (the image's OpenCV), glibc's srand/rand and libstdc++'s std::sort:
                       cv::resize(INTER_LINEAR) chain, copyMakeBorder(REFLECT_101), GaussianBlur(7x7, 2),
                       cv::FAST(th, nonmax), cv::fastAtan2, BFMatcher(HAMMING).knnMatch(k=2), rand() streams,
                       std::sort permutations of DMatch-like records.  These pin the oracle's restatements.
  pipeline_oracle.npz  outputs of the oracle itself (ORB keypoints/descriptors, Matcher::match, RansacSE3) on a small
                       seeded 320x240 three-frame sequence.  The reference (PCL/Eigen/OpenCV C++ build) cannot be
                       compiled in this image, so these are regression vectors of the restatement, NOT reference
                       outputs ("parity unpinned" for the composed pipeline).

Neither the tests nor the product read /root/reference; nothing here does either.
"""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

SORT_SRC = r"""
#include <algorithm>
#include <cstdio>
#include <vector>
struct M { float d; int i; bool operator<(const M& o) const { return d < o.d; } };
int main() {
    int n;
    while (scanf("%d", &n) == 1) {
        std::vector<M> v(n);
        for (int k = 0; k < n; k++) { if (scanf("%f", &v[k].d) != 1) return 1; v[k].i = k; }
        std::sort(v.begin(), v.end());
        for (int k = 0; k < n; k++) printf("%d ", v[k].i);
        printf("\n");
    }
}
"""


def std_sort_permutations(key_lists):
    """Permutation libstdc++'s std::sort applies to records ordered by `distance` only (cv::DMatch::operator<)."""
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, "s.cpp"), os.path.join(d, "s")
        open(src, "w").write(SORT_SRC)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, src])
        inp = "".join(f"{len(k)} " + " ".join(str(float(x)) for x in k) + "\n" for k in key_lists)
        out = subprocess.run([exe], input=inp, capture_output=True, text=True, check=True).stdout
    return [np.array(line.split(), np.int32) for line in out.strip().split("\n")]


def smooth_image(rng, w, h):
    import cv2

    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    img = cv2.GaussianBlur(img, (5, 5), 1.2)
    # blocks with edges and corners so that FAST fires at both thresholds
    for _ in range(40):
        x, y = int(rng.integers(0, w - 12)), int(rng.integers(0, h - 12))
        img[y:y + int(rng.integers(4, 12)), x:x + int(rng.integers(4, 12))] = int(rng.integers(0, 256))
    return img


def make_primitives(path):
    import cv2

    rng = np.random.default_rng(20241104)
    out = {"cv2_version": np.array(cv2.__version__)}
    # cv::resize chain (the extractor's pyramid sizes for 320x240, scale 1.2) + REFLECT_101 ring + 7x7 blur
    img = smooth_image(rng, 320, 240)
    out["img"] = img
    # level sizes exactly as ORBextractor computes them: cvRound(w * invScaleFactor[l]) with float scale factors
    sf = [np.float32(1.0)]
    for l in range(1, 8):
        sf.append(np.float32(sf[-1] * np.float32(1.2)))
    sizes = [(int(np.rint(np.float32(320) * (np.float32(1.0) / s))), int(np.rint(np.float32(240) * (np.float32(1.0) / s)))) for s in sf]
    out["level_sizes"] = np.array(sizes, np.int32)
    prev = img
    for l, (w, h) in enumerate(sizes):
        cur = prev if l == 0 else cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
        out[f"level{l}"] = cur
        out[f"blur{l}"] = cv2.GaussianBlur(cur, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        prev = cur
    out["border19_level3"] = cv2.copyMakeBorder(out["level3"], 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    # cv::FAST on whole small images, both thresholds, with and without non-max suppression
    for k, (w, h) in enumerate([(96, 72), (37, 61), (64, 64)]):
        im = smooth_image(rng, w, h)
        out[f"fast_img{k}"] = im
        for th in (7, 20):
            for nm in (0, 1):
                kp = cv2.FastFeatureDetector_create(th, bool(nm)).detect(im, None)
                out[f"fast{k}_th{th}_nm{nm}"] = np.array([(p.pt[0], p.pt[1], p.response) for p in kp], np.float32).reshape(-1, 3)
    # cv::fastAtan2 (degrees)
    y = rng.integers(-200000, 200000, 4096).astype(np.float32)
    x = rng.integers(-200000, 200000, 4096).astype(np.float32)
    y[:8] = 0
    x[4:12] = 0
    out["atan2_y"], out["atan2_x"] = y, x
    out["atan2_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    # BFMatcher(NORM_HAMMING).knnMatch(k=2) with exact ties
    q = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    t[10] = t[20] = t[5]
    q[5] = t[10]
    t[100:110] = t[100]
    q[7] = t[100]
    kn = cv2.BFMatcher_create(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    out["knn_q"], out["knn_t"] = q, t
    out["knn_idx"] = np.array([[m[0].trainIdx, m[1].trainIdx] for m in kn], np.int32)
    out["knn_dist"] = np.array([[m[0].distance, m[1].distance] for m in kn], np.float32)
    # glibc srand/rand
    libc = C.CDLL("libc.so.6")
    for seed in (42, 1, 0, 123456789):
        libc.srand(seed)
        out[f"rand_seed{seed}"] = np.array([libc.rand() for _ in range(2000)], np.int32)
    # libstdc++ std::sort on integer-valued float keys with many ties (DMatch distances are Hamming counts)
    keys = [rng.integers(0, 100, n).astype(np.float32) for n in (3, 16, 17, 40, 333, 1000, 2000)]
    keys.append(np.zeros(500, np.float32))
    keys.append(np.arange(700, 0, -1).astype(np.float32))
    perms = std_sort_permutations(keys)
    for k, (a, p) in enumerate(zip(keys, perms)):
        out[f"sort_keys{k}"], out[f"sort_perm{k}"] = a, p
    out["sort_cases"] = np.array(len(keys))
    np.savez_compressed(path, **out)
    return out


def make_pipeline(path):
    """Oracle outputs for a 320x240 crop of the synthetic TUM-fr1-shaped sequence (three frames, two tracked pairs)."""
    import orclib
    import synth

    cam, frames = synth.sequence("tum_fr1", n=3, seed=4321)
    x0, y0, w, h = 160, 120, 320, 240
    cam_c = dict(fx=cam["fx"], fy=cam["fy"], cx=cam["cx"] - x0, cy=cam["cy"] - y0, factor=5000.0)
    out = {"cam": np.array([cam_c["fx"], cam_c["fy"], cam_c["cx"], cam_c["cy"], cam_c["factor"]], np.float64)}
    nfeat = 500
    orb = orclib.Orb(nfeat)
    rng, latch = orclib.Rng(42), orclib.Latch()
    prev = None
    for i, (g, d) in enumerate(frames):
        gray = np.ascontiguousarray(g[y0:y0 + h, x0:x0 + w])
        d16 = np.rint(np.nan_to_num(d[y0:y0 + h, x0:x0 + w], nan=0.0) * 5000.0).clip(0, 65535).astype(np.uint16)
        depth = d16.astype(np.float32) * np.float32(1.0 / 5000.0)  # cv::Mat::convertTo(CV_32F, 1/factor) of the 16-bit PNG
        out[f"gray{i}"], out[f"depth16_{i}"] = gray, d16
        k, de = orb.extract(gray)
        p3 = synth.unproject(cam_c, np.stack([k["x"], k["y"]], 1), depth)
        for f in ("x", "y", "size", "angle", "response", "octave"):
            out[f"kp{i}_{f}"] = k[f]
        out[f"desc{i}"], out[f"xyz{i}"] = de, p3
        outl = np.zeros(len(k), np.uint8)
        if prev is not None:
            m = orclib.match(prev[1], de, prev[3], prev[2], p3, 0.9, True)
            r = orclib.ransac_se3(prev[2], p3, m, rng, latch, 200, 10, 3.0, 4)
            if len(m) >= 10:
                outl[m["trainIdx"]] = 1
            if r["ok"]:
                outl[r["inliers"]["trainIdx"]] = 0
            out[f"match{i}_q"], out[f"match{i}_t"], out[f"match{i}_d"] = m["queryIdx"], m["trainIdx"], m["distance"]
            out[f"se3_{i}_ok"] = np.array(int(r["ok"]))
            out[f"se3_{i}_T"], out[f"se3_{i}_rmse"] = r["T"], np.array(r["rmse"], np.float32)
            out[f"se3_{i}_inl_t"] = r["inliers"]["trainIdx"]
            out[f"se3_{i}_iters"] = np.array(r["iters_executed"])
        out[f"outlier{i}"] = outl
        prev = (k, de, p3, outl)
    out["rng_next_after"] = np.array(rng.next())
    out["nfeatures"] = np.array(nfeat)
    np.savez_compressed(path, **out)
    return out


TUM_FR1 = dict(fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989,
               dist=[0.262383, -0.953104, -0.005358, 0.002628, 1.163314], factor=5000.0)


def make_frame_prep(path):
    """cv2 outputs for Frame's preprocessing (row N1): cvtColor(BGR2GRAY) and undistortPoints with the TUM fr1 camera."""
    import cv2

    rng = np.random.default_rng(20241105)
    bgr = np.stack([smooth_image(rng, 320, 240) for _ in range(3)], 2)
    out = {"bgr": bgr, "gray": cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)}
    c = TUM_FR1
    K = np.array([[c["fx"], 0, c["cx"]], [0, c["fy"], c["cy"]], [0, 0, 1]], np.float32)
    D = np.array(c["dist"], np.float32).reshape(5, 1)
    pts = np.stack([rng.uniform(0, 640, 3000), rng.uniform(0, 480, 3000)], 1).astype(np.float32)
    pts[:4] = [[0, 0], [640, 0], [0, 480], [640, 480]]           # the corners computeImageBounds undistorts
    pts[4:1004] = np.rint(pts[4:1004] * 1.2) / np.float32(1.2)   # keypoint-like positions (integer * scale factor)
    out["pts"] = pts
    out["pts_un"] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    out["cam"] = np.array([c["fx"], c["fy"], c["cx"], c["cy"]] + c["dist"] + [c["factor"]], np.float64)
    np.savez_compressed(path, **out)
    return out


def make_projection(path):
    """cv2 outputs for the primitives of Matcher::projectionMatch (row N2): the 3x3 * 3x1 + 3x1 cv::Mat expression
    (cv::gemm) of Frame::unprojectWorld / Matcher.cpp:52 and the Hamming norm of descriptorDistance."""
    import cv2

    rng = np.random.default_rng(20241106)
    n = 4000
    A = rng.normal(0, 1, (n, 3, 3)).astype(np.float32)
    A[: n // 2] = np.stack([cv2.Rodrigues(rng.normal(0, 0.4, 3))[0] for _ in range(n // 2)]).astype(np.float32)  # rotations
    b = (rng.normal(0, 2, (n, 3)) * rng.choice([1e-3, 1, 50], (n, 1))).astype(np.float32)
    c = rng.normal(0, 1, (n, 3)).astype(np.float32)
    g = np.stack([cv2.gemm(A[i], b[i].reshape(3, 1), 1.0, c[i].reshape(3, 1), 1.0).reshape(3) for i in range(n)])
    da = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    db[:100] = da[:100] ^ (rng.random((100, 32)) < 0.1).astype(np.uint8)
    ham = np.array([cv2.norm(da[i], db[i], cv2.NORM_HAMMING) for i in range(len(da))], np.float64)
    out = {"A": A, "b": b, "c": c, "gemm": g.astype(np.float32), "desc_a": da, "desc_b": db, "hamming": ham}
    np.savez_compressed(path, **out)
    return out


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "projection":  # the other fixtures stay as committed
        make_projection(os.path.join(HERE, "projection_cv2.npz"))
        print("projection_cv2.npz", os.path.getsize(os.path.join(HERE, "projection_cv2.npz")), "bytes")
        sys.exit(0)
    make_projection(os.path.join(HERE, "projection_cv2.npz"))
    make_frame_prep(os.path.join(HERE, "frame_prep_cv2.npz"))
    print("frame_prep_cv2.npz", os.path.getsize(os.path.join(HERE, "frame_prep_cv2.npz")), "bytes")
    a = make_primitives(os.path.join(HERE, "cv2_primitives.npz"))
    b = make_pipeline(os.path.join(HERE, "pipeline_oracle.npz"))
    for n in ("cv2_primitives.npz", "pipeline_oracle.npz"):
        print(n, os.path.getsize(os.path.join(HERE, n)), "bytes")
    print("pairs:", [(int(b[f"se3_{i}_ok"]), len(b[f"match{i}_q"]), len(b[f"se3_{i}_inl_t"]), int(b[f"se3_{i}_iters"])) for i in (1, 2)])
