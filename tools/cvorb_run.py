"""Runs the cv::ORB branch alone (one context, one 64-frame batch per step, device-resident frames) and prints the
per-kernel times from the library's CUDA-event profile; the target of the ncu captures of the N3 kernels."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")]
import rgbdfe  # noqa: E402
import synth  # noqa: E402
import torch  # noqa: E402

W, H = 640, 480
B = int(os.environ.get("CVORB_BATCH", "64"))
NFEAT = int(os.environ.get("CVORB_NFEAT", "2000"))
STEPS = int(os.environ.get("CVORB_STEPS", "6"))
cam, frames = synth.sequence("icl", n=8, seed=2345, noise_sigma=0.0, invalid_frac=0.0)
gray = torch.from_numpy(np.stack([frames[i % 8][0] for i in range(B)])).cuda()
depth = torch.from_numpy(np.stack([(frames[i % 8][1] * 5000).astype(np.uint16) for i in range(B)])).cuda()
L = rgbdfe.lib()
ctx = rgbdfe.Context(0, W, H, NFEAT, B)
ctx.cvorb_configure(NFEAT, 1.2, 8)
fr = [rgbdfe.Frame(ctx) for _ in range(B)]
fa = (C.c_void_p * B)(*[f.h for f in fr])
gp = (C.c_void_p * B)(*[gray[i].data_ptr() for i in range(B)])
dp = (C.c_void_p * B)(*[depth[i].data_ptr() for i in range(B)])


def step():
    ctx._chk(L.rgbdfe_frames_extract_raw(ctx.h, fa, B, gp, W, H, W, dp, W, C.c_float(1 / 5000.0), C.c_float(cam["fx"]), C.c_float(cam["fy"]),
                                         C.c_float(cam["cx"]), C.c_float(cam["cy"])))


for _ in range(3):
    step()
ctx.sync()
t0 = time.perf_counter()
for _ in range(STEPS):
    step()
ctx.sync()
dt = time.perf_counter() - t0
print(f"cvorb alone: {STEPS * B / dt:.0f} frames/s ({dt / STEPS * 1e3:.3f} ms per {B}-frame batch), {len(fr[0].download()[0])} keypoints")
ctx._chk(L.rgbdfe_profile_enable(ctx.h, 1))
for _ in range(3):
    step()
a_ms, a_cnt = (C.c_double * 16)(), (C.c_longlong * 16)()
ctx._chk(L.rgbdfe_profile_read(ctx.h, a_ms, a_cnt, 16))
names = ["pyramid(8 launches)", "k_cv_fast", "k_cv_blur", "k_cv_select+gather", "k_cv_orient_describe"]
for i, n in enumerate(names):
    print(f"  {n:28s} {a_ms[i] / max(a_cnt[i], 1):.4f} ms")
