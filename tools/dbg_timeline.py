"""Developer aid: phase timeline (device %globaltimer) of the match/RANSAC chain for one pair of the bench workload."""
import ctypes as C
import sys
sys.path[:0] = ["tests", "tools", "rgbd-slam_b200/py", "."]
import numpy as np, rgbdfe, synth, bench

nfeat = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
cam, frames = bench.make_sequence(6)
ctx = rgbdfe.Context(0, 640, 480, nfeat, 8)
ctx.orb_configure(nfeat, 1.2, 8, 20, 7)
ctx.rng_seed(42)
fr = [rgbdfe.Frame(ctx) for _ in frames]
ctx.frames_extract(fr, [f[0] for f in frames], [f[1] for f in frames], cam)
L = rgbdfe.lib()
for rep in range(3):
    for i in range(1, len(frames)):
        L.rgbdfe_debug_timeline(ctx.h, 1, None, 0)
        ctx.track(fr[i - 1], fr[i], 0, 0.9, True, 500, 10, 3.0, 4, True)
        res = ctx.track_results(0, 1)[0]
        t = np.zeros(1024, np.int64)
        L.rgbdfe_debug_timeline(ctx.h, 0, t.ctypes.data_as(C.c_void_p), 1024)
        if rep < 2:
            continue
        t0 = t[0]
        prep = [(int(t[k] - t0) / 1e3 if t[k] else None) for k in range(5)]
        print(f"pair {i}: M={res['n_matches']} inl={res['n_inliers']} iters={res['iters_executed']}  prepare us: filter_end={prep[1]} sort_end={prep[2]} sync={prep[3]} gather_end={prep[4]}")
        lv = []
        for q in range(40):
            if not t[400 + 2 * q]:
                break
            lv.append(f"{(t[400 + 2 * q] - t0) / 1e3:.1f}:{t[401 + 2 * q]}")
        print("   sort levels (us:ranges) " + " ".join(lv))
        for q in range(1, 12):
            b = 480 + 8 * q
            if not t[b]:
                continue
            print(f"     lvl{q-1} range0 n={t[b+6]}: median={(t[b+1]-t[b])/1e3:.2f} LR={(t[b+2]-t[b+1])/1e3:.2f} K={(t[b+3]-t[b+2])/1e3:.2f} swap={(t[b+4]-t[b+3])/1e3:.2f} enqueue={(t[b+5]-t[b+4])/1e3:.2f}")
        h0 = t[16]
        line = [f"hyp0 start +{(h0 - t0) / 1e3:.1f}us fit0={(t[17] - h0) / 1e3:.1f}"]
        for r in range(1, 20):
            b = 16 + r * 8
            if not t[b]:
                break
            sc = (t[b + 1] - t[b]) / 1e3 if t[b + 1] else -1
            fl = (t[b + 2] - t[b + 1]) / 1e3 if t[b + 2] else -1
            ft = (t[b + 3] - t[b + 2]) / 1e3 if t[b + 3] else -1
            fx = ""
            if t[b + 4]:
                fx = f" (prefix={(t[b + 4] - t[b + 2]) / 1e3:.1f} div={(t[b + 5] - t[b + 4]) / 1e3:.1f} chain={(t[b + 6] - t[b + 5]) / 1e3:.1f} svd={(t[b + 3] - t[b + 6]) / 1e3:.1f})"
            line.append(f"r{r}[n={t[b + 7]} score={sc:.1f} scanfill={fl:.1f} sum|fit={ft:.1f}{fx}]")
        if t[200]:
            line.append(f"sm_clock={(t[301] - t[300]) / max(t[200] - t[16], 1) * 1e3:.0f}MHz")
            line.append(f"publish@{(t[200] - h0) / 1e3:.1f} replay_done@{(t[201] - h0) / 1e3:.1f}")
        print("   " + " ".join(line))
ctx.close()
