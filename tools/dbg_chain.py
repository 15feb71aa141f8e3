import sys, os
sys.path[:0] = ["tests", "tools", "rgbd-slam_b200/py"]
import numpy as np, rgbdfe, synth
cam, frames = synth.sequence("tum_fr1", n=6, seed=1234)
out = {}
for chain in (False, True):
    ctx = rgbdfe.Context(0, 640, 480, 1000, 8)
    ctx.rng_seed(42)
    fr = [rgbdfe.Frame(ctx) for _ in frames]
    ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
    if chain:
        ctx.track_sequence(fr, 0)
    else:
        for i in range(1, len(frames)):
            ctx.track(fr[i - 1], fr[i], i - 1)
    res = ctx.track_results(0, len(frames) - 1)
    print(chain, [(int(r["n_matches"]), int(r["n_inliers"]), int(r["iters_executed"]), int(r["ok"])) for r in res])
    print("  flags", [int(f.download()[3].sum()) for f in fr], "rng idx", ctx.rng_get_state()[34])
    ctx.close()
