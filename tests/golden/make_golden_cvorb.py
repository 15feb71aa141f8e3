#!/usr/bin/env python
"""Generates tests/golden/cvorb_cv2.npz (run once in the build container; the output is committed).

Row N3: outputs of the REAL OpenCV (cv2 4.13.0 of this image) for the Extractor(ORB, ORB) call sequence of the
reference (Features/Extractor.cpp:55-59,163-166,216-218):
    det = cv2.ORB_create(nfeatures, 1.2, 8); des = cv2.ORB_create()
    kps = det.detect(gray); kps, desc = des.compute(gray, kps)
plus the two OpenCV primitives the branch adds to the path: resize(INTER_LINEAR_EXACT) and the float sepFilter2D
blur that cv::ORB's GaussianBlur call resolves to on a pyramid submatrix.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")]
import synth  # noqa: E402

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def kparr(k):
    return np.array([(p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave, p.class_id) for p in k], dtype=KP)


def extractor_orb_orb(gray, nfeatures):
    det = cv2.ORB_create(nfeatures, 1.2, 8)
    des = cv2.ORB_create()
    k = det.detect(gray, None)
    if len(k) > nfeatures:  # Extractor.cpp:56-57 (cv2 has no retainBest binding; cv::ORB never returns more unless ties)
        raise SystemExit("fixture needs a case without overshoot")
    k, d = des.compute(gray, k)
    return kparr(k), d


def main():
    out = {}
    cam, frames = synth.sequence("tum_fr1", n=1, seed=21)
    g = cv2.resize(frames[0][0], (320, 240), interpolation=cv2.INTER_AREA)
    out["img_a"] = g
    out["kps_a"], out["desc_a"] = extractor_orb_orb(g, 500)
    # a periodic image: every FAST and Harris response repeats, so both retainBest calls cut through ties
    rng = np.random.default_rng(77)
    patch = cv2.GaussianBlur(rng.integers(0, 256, (48, 64), dtype=np.uint8), (5, 5), 1.2)
    t = np.tile(patch, (5, 5))[:232, :312].copy()
    out["img_b"] = t
    out["kps_b"], out["desc_b"] = extractor_orb_orb(t, 300)
    # primitives
    src = rng.integers(0, 256, (97, 131), dtype=np.uint8)
    out["resize_src"] = src
    out["resize_dst"] = cv2.resize(src, (109, 81), interpolation=cv2.INTER_LINEAR_EXACT)
    kx = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    bsrc = cv2.GaussianBlur(rng.integers(0, 256, (70, 107), dtype=np.uint8), (5, 5), 1.5)
    out["blur_src"] = bsrc
    out["blur_dst"] = cv2.sepFilter2D(bsrc, cv2.CV_8U, kx, kx, borderType=cv2.BORDER_REFLECT_101)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(HERE, "cvorb_cv2.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()})


if __name__ == "__main__":
    main()
