#!/usr/bin/env python
"""Per-kernel times of the verification pass (kNN / prepare+filter / hypotheses) for the three kNN kernels, CUDA events
through rgbdfe_profile_*: one query against n_kf keyframes of n_feat features."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")):
    sys.path.insert(0, p)
import numpy as np
import rgbdfe
import synth


def run(n_kf, n_feat, umma, reps=5, dbg=0):
    qd, qp, kd, kp, _ = synth.synth_pair_sets(n_kf, n_feat, seed=5678, overlap_frac=0.05)
    ctx = rgbdfe.Context(0, 640, 480, n_feat, 1)
    ctx.depth_cov_set(1, (0.01 * 2.0 * 2.0) ** 2)
    ctx.set_option("knn_umma", umma)
    ctx.set_option("knn_dbg", dbg)
    db = ctx.kfdb_create(kd, kp)
    L = rgbdfe.lib()
    L.rgbdfe_profile_name.restype = C.c_char_p
    for _ in range(2):
        ctx.verify_batch(db, n_kf, qd, qp)
    ctx._chk(L.rgbdfe_profile_enable(ctx.h, 1))
    for _ in range(reps):
        ctx.verify_batch(db, n_kf, qd, qp)
    ms, cnt = (C.c_double * 16)(), (C.c_longlong * 16)()
    ctx._chk(L.rgbdfe_profile_read(ctx.h, ms, cnt, 16))
    ctx._chk(L.rgbdfe_profile_enable(ctx.h, 0))
    out = {L.rgbdfe_profile_name(i).decode(): round(ms[i] / max(cnt[i], 1), 4) for i in range(10) if cnt[i]}
    ctx.kfdb_destroy(db)
    ctx.close()
    return out


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "dbg":
    for dbg in (0, 1, 2, 4, 3, 5, 6, 7):
        print("dbg", dbg, run(64, 2000, 2, dbg=dbg)["k_knn2_umma"], run(4096, 1000, 2, dbg=dbg)["k_knn2_umma"])
    sys.exit(0)
if __name__ == "__main__":
    for n_kf, n_feat in ((4096, 1000), (64, 2000), (512, 1000)):
        for umma in (2, 1, 0):
            print(n_kf, n_feat, "knn_umma", umma, run(n_kf, n_feat, umma))
