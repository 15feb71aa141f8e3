"""Multi-GPU path of the loop-pair verification, exercised on CPU: two gloo ranks shard the keyframes, verify their
block (the CPU oracle stands in for rgbdfe_verify_batch here; only the host logic is under test) and all_gather the
84-byte results; the gathered list must equal the single-process list, in keyframe order."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_KF, N_FEAT, SEED = 7, 200, 900  # 7 keyframes over 2 ranks: blocks of 4 and 3 (ragged)


def _oracle_verify(kd, kp):
    import orclib
    import rgbdfe

    def run(lo, hi, qd, qp, seed):
        out = np.zeros(hi - lo, rgbdfe.SE3_DTYPE)
        for k in range(lo, hi):
            m = orclib.match(kd[k], qd, np.zeros(N_FEAT, np.uint8), kp[k], qp, 0.9, True)
            out[k - lo]["n_matches"] = len(m)
            out[k - lo]["rmse"] = 1e6
            out[k - lo]["T21"] = np.eye(4, dtype=np.float32).ravel()
            if len(m) < 30:
                continue
            rng, latch = orclib.Rng(seed + k), orclib.Latch()
            latch.latched.value, latch.cov.value = 1, (0.01 * 2.0 * 2.0) ** 2
            r = orclib.ransac_se3(kp[k], qp, m, rng, latch, 200, 30, 3.0, 4)
            out[k - lo]["ok"], out[k - lo]["rmse"], out[k - lo]["T21"] = int(r["ok"]), r["rmse"], r["T"].ravel()
            out[k - lo]["n_inliers"], out[k - lo]["iters_executed"] = len(r["inliers"]), r["iters_executed"]
        return out

    return run


def _worker(rank, world, port, q):
    import torch.distributed as dist

    import loopverify
    import rgbdfe
    import synth

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    qd, qp, kd, kp, _ = synth.synth_pair_sets(N_KF, N_FEAT, seed=31, overlap_frac=0.7)
    v = loopverify.ShardedVerifier(N_KF, rank, world, _oracle_verify(kd, kp))
    res = v.verify(qd, qp, SEED, rgbdfe.SE3_DTYPE)
    # batched relocalisation (configs[4] shape): 3 queries x N_KF keyframes, one all_gather for the whole batch
    qs = [qd, np.roll(qd, 5, axis=0), qd[::-1].copy()]
    ps = [qp, np.roll(qp, 5, axis=0), qp[::-1].copy()]
    many = v.verify_many(qs, ps, SEED, rgbdfe.SE3_DTYPE)
    q.put((rank, res.tobytes(), (v.lo, v.hi), many.tobytes()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_cover_and_are_ragged_safe():
    import loopverify

    for n in (1, 7, 8, 4096, 4099):
        for w in (1, 2, 3, 8):
            r = [loopverify.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_two_rank_gloo_gather_equals_single_process():
    import torch.multiprocessing as mp

    import loopverify
    import rgbdfe
    import synth

    qd, qp, kd, kp, _ = synth.synth_pair_sets(N_KF, N_FEAT, seed=31, overlap_frac=0.7)
    single = _oracle_verify(kd, kp)(0, N_KF, qd, qp, SEED)
    assert single["ok"].sum() >= 2  # the fixture has real loop closures
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = [q.get(timeout=120) for _ in procs]
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    blocks = sorted(g[2] for g in got)
    assert blocks == [(0, 4), (4, 7)]
    run = _oracle_verify(kd, kp)
    qs = [qd, np.roll(qd, 5, axis=0), qd[::-1].copy()]
    ps = [qp, np.roll(qp, 5, axis=0), qp[::-1].copy()]
    many_single = np.stack([run(0, N_KF, qs[i], ps[i], SEED + i * N_KF) for i in range(3)])
    assert many_single[1]["ok"].sum() >= 2  # a permuted query still closes its loops
    for _, raw, _, many in got:  # every rank holds the full list, in keyframe order
        res = np.frombuffer(raw, dtype=rgbdfe.SE3_DTYPE)
        assert len(res) == N_KF and res.tobytes() == single.tobytes()
        assert many == many_single.tobytes()  # [3, N_KF] records, independent of the sharding
    assert loopverify.first_success(single) == int(np.nonzero(single["ok"])[0][0])
