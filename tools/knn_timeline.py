#!/usr/bin/env python
"""clock64 stamps of the six hand-offs of k_knn2_umma_ws for the first tiles of CTA 0 (developer aid)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")):
    sys.path.insert(0, p)
import numpy as np
import rgbdfe
import synth

dbg = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n_kf, n_feat = 64, 2000
qd, qp, kd, kp, _ = synth.synth_pair_sets(n_kf, n_feat, seed=5678, overlap_frac=0.05)
ctx = rgbdfe.Context(0, 640, 480, n_feat, 1)
ctx.depth_cov_set(1, (0.01 * 2.0 * 2.0) ** 2)
ctx.set_option("knn_dbg", dbg)
db = ctx.kfdb_create(kd, kp)
L = rgbdfe.lib()
for _ in range(2):
    ctx.verify_batch(db, n_kf, qd, qp)
buf = (C.c_longlong * 1024)()
ctx._chk(L.rgbdfe_debug_timeline(ctx.h, 1, None, 0))
ctx.verify_batch(db, n_kf, qd, qp)
ctx._chk(L.rgbdfe_debug_timeline(ctx.h, 0, buf, 1024))
t = np.array(buf[:384], dtype=np.int64).reshape(6, 64)
t0 = t[t > 0].min()
names = ["E:acc_empty ok", "E:b_full arrive", "I:b_full ok", "I:commit", "D:acc_full ok", "D:folded"]
print("seq  " + "  ".join(f"{n:>16s}" for n in names))
for s in range(40):
    print(f"{s:3d}  " + "  ".join(f"{(t[e, s] - t0) if t[e, s] > 0 else -1:16d}" for e in range(6)))
mm = np.array(buf[400:432], dtype=np.int64).reshape(4, 8)
print("MMA issue stamps of tiles 20..23 (clocks after the tile's first):")
for r in mm:
    print("   ", list((r - r[0]).astype(int)))
d = np.diff(t[5, 8:40])
print("steady-state clocks per tile (drain):", d.mean(), d.min(), d.max())
