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


def test_c_shard_rule_and_unpacking_match_the_host_rule():
    """rgbdfe_shard_range / rgbdfe_gathered_to_keyframe_order (host arithmetic of the C ABI, no GPU needed)."""
    import loopverify
    import rgbdfe

    for n in (1, 7, 8, 4096, 4099):
        for w in (1, 2, 3, 8):
            for r in range(w):
                assert rgbdfe.shard_range(n, r, w) == loopverify.shard_range(n, r, w)
    n_kf, world, Q = 11, 3, 2
    per = -(-n_kf // world)
    full = np.zeros((Q, n_kf), rgbdfe.SE3_DTYPE)
    full["n_matches"] = np.arange(Q * n_kf).reshape(Q, n_kf) + 1
    g = np.zeros((world, Q, per), rgbdfe.SE3_DTYPE)
    for r in range(world):
        lo, hi = loopverify.shard_range(n_kf, r, world)
        g[r, :, : hi - lo] = full[:, lo:hi]
    assert rgbdfe.gathered_to_keyframe_order(g.reshape(-1), world, Q, n_kf).tobytes() == full.tobytes()


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


# ---- on hardware: the C-ABI path itself (rgbdfe_verify_sharded_*), 1 rank and 2 ranks over NCCL ----
G_KF, G_FEAT = 37, 400  # 37 keyframes over 2 ranks: blocks of 19 and 18


def _gpu_queries():
    import synth

    qd, qp, kd, kp, _ = synth.synth_pair_sets(G_KF, G_FEAT, seed=91, overlap_frac=0.5)
    qs = np.stack([qd, np.roll(qd, 7, axis=0), qd[::-1]]).copy()
    ps = np.stack([qp, np.roll(qp, 7, axis=0), qp[::-1]]).copy()
    return qs, ps, kd, kp


@pytest.mark.gpu
def test_verify_sharded_one_rank_equals_verify_batch_and_pipelines():
    import loopverify
    import rgbdfe

    qs, ps, kd, kp = _gpu_queries()
    v = loopverify.CudaShardedVerifier(0, 0, 1, G_KF, kd, kp, max_feats=G_FEAT)
    try:
        many = v.verify(qs, ps, seed=900)  # 3 queries, one call
        assert many.shape == (3, G_KF) and many["ok"].sum() >= 6
        for q in range(3):
            one = v.ctx.verify_batch(v.db, G_KF, qs[q], ps[q], 0.9, 30, 200, 30, 3.0, 4, seed=900, first_pair_index=q * G_KF)
            assert one.tobytes() == many[q].tobytes()
        # pipelined single queries on rotating slots: same bytes as the blocking calls, whatever is in flight
        for q in range(3):
            v.enqueue(qs[q], ps[q], G_FEAT, 1, q, seed=900 + q * G_KF)
        for q in range(3):
            assert v.collect(q, 1)[0].tobytes() == many[q].tobytes()
        with pytest.raises(rgbdfe.RgbdfeError):  # a slot must be collected before it is reused
            v.enqueue(qs[0], ps[0], G_FEAT, 1, 0)
            v.enqueue(qs[0], ps[0], G_FEAT, 1, 0)
        v.collect(0, 1)
        # an unlatched context refuses batched verification
        c2 = rgbdfe.Context(0, 640, 480, G_FEAT, 1)
        db2 = c2.kfdb_create(kd, kp)
        with pytest.raises(rgbdfe.RgbdfeError, match="latch"):
            c2.verify_batch(db2, G_KF, qs[0], ps[0])
        c2.kfdb_destroy(db2)
        c2.close()
    finally:
        v.close()


def _nccl_worker(rank, world, id_q, out_q):
    import loopverify
    import rgbdfe

    try:
        if rank == 0:
            ident = rgbdfe.nccl_unique_id()
            for _ in range(world - 1):
                id_q.put(ident)
        else:
            ident = id_q.get(timeout=120)
        qs, ps, kd, kp = _gpu_queries()
        lo, hi = rgbdfe.shard_range(G_KF, rank, world)
        v = loopverify.CudaShardedVerifier(rank, rank, world, G_KF, kd[lo:hi], kp[lo:hi], ident, max_feats=G_FEAT)
        many = v.verify(qs, ps, seed=900)
        single = [v.verify(qs[q], ps[q], seed=900 + q * G_KF) for q in range(3)]
        out_q.put((rank, many.tobytes(), b"".join(s.tobytes() for s in single)))
        v.close()
    except Exception as ex:  # surface the failure in the parent
        out_q.put((rank, repr(ex), b""))


@pytest.mark.gpu
def test_verify_sharded_two_gpus_over_nccl_equals_one_gpu():
    import torch
    import torch.multiprocessing as mp

    import loopverify

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    qs, ps, kd, kp = _gpu_queries()
    v = loopverify.CudaShardedVerifier(0, 0, 1, G_KF, kd, kp, max_feats=G_FEAT)
    want = v.verify(qs, ps, seed=900).tobytes()
    v.close()
    ctx = mp.get_context("spawn")
    id_q, out_q = ctx.Queue(), ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, id_q, out_q)) for r in range(2)]
    [p.start() for p in procs]
    got = [out_q.get(timeout=300) for _ in procs]
    [p.join(timeout=60) for p in procs]
    for rank, many, single in got:
        assert isinstance(many, bytes), f"rank {rank}: {many}"
        assert many == want and single == want  # every rank holds the full list, identical to the 1-GPU result
