"""Sharded loop-pair verification (PoseGraph::createLocalEdges / detectLoop, Solver/PoseGraph.cpp:128-155,245-287).

One process per GPU.  The keyframe database is block-sharded by keyframe index and resident on its rank;
a query frame is handed to every rank, each rank runs match + RansacSE3 for its block through
rgbdfe_verify_batch, and ONE collective (all_gather of the 84-byte per-pair results) rebuilds the list in
keyframe order.  Per-pair rand() streams are srand(seed + global pair index), so the result does not depend
on the number of ranks.  The data path has no other exchange step.

Host logic only: the compute is the CUDA library's; `local_verify` may be any callable with the same contract,
which is how the CPU (gloo) tests drive this module without a GPU.
"""
import numpy as np

RESULT_BYTES = 84  # sizeof(rgbdfe_se3_result)


def shard_range(n_kf, rank, world):
    """Contiguous block of keyframes owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_kf, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def first_success(results, min_index=0):
    """detectLoop takes the FIRST candidate, in candidate order, whose RansacSE3 succeeds (PoseGraph.cpp:283)."""
    ok = np.nonzero(results["ok"][min_index:] != 0)[0]
    return int(ok[0]) + min_index if len(ok) else -1


class ShardedVerifier:
    """Host logic of the sharded path with a pluggable per-block verifier (the CPU gloo tests drive it with the oracle):
    `local_verify(kf_lo, kf_hi, q_desc, q_xyz, seed)` returns the structured results of keyframes [kf_lo, kf_hi) with
    per-pair stream srand(seed + global index).  The exchanged layout and its unpacking are the C ABI's
    ([rank][query][ceil(n_kf / world)] records -> rgbdfe_gathered_to_keyframe_order)."""

    def __init__(self, n_kf, rank, world, local_verify, device="cpu"):
        self.n_kf, self.rank, self.world, self.device = n_kf, rank, world, device
        self.lo, self.hi = shard_range(n_kf, rank, world)
        self.local_verify = local_verify

    def verify_many(self, q_descs, q_xyzs, seed, dtype):
        """Batched relocalisation (BASELINE configs[4]: Q query frames x n_kf keyframes, all pairs): every rank verifies
        all Q queries against its keyframe block, then ONE all_gather moves the Q x block results.  Pair (q, k) uses the
        stream srand(seed + q * n_kf + k), so the result is independent of the sharding.  Returns [Q, n_kf] records."""
        import torch

        import rgbdfe

        Q = len(q_descs)
        per = -(-self.n_kf // self.world)
        block = np.zeros((Q, per), dtype)
        for q in range(Q):
            block[q, : self.hi - self.lo] = self.local_verify(self.lo, self.hi, q_descs[q], q_xyzs[q], seed + q * self.n_kf)
        buf = torch.from_numpy(block.view(np.uint8).reshape(-1).copy()).to(self.device)
        if self.world == 1:
            out = buf
        else:
            import torch.distributed as dist

            out = torch.empty(self.world * buf.numel(), dtype=torch.uint8, device=buf.device)
            dist.all_gather_into_tensor(out, buf)
        gathered = np.frombuffer(out.cpu().numpy().tobytes(), dtype=dtype)
        return rgbdfe.gathered_to_keyframe_order(gathered, self.world, Q, self.n_kf)

    def verify(self, q_desc, q_xyz, seed, dtype):
        return self.verify_many([q_desc], [q_xyz], seed, dtype)[0]


class CudaShardedVerifier:
    """One rank of the sharded verification through the C ABI only (what a C++ PoseGraph would call): this rank's
    keyframe block resident in a rgbdfe_kfdb, an NCCL communicator made by rgbdfe_nccl_comm_create from a 128-byte id
    that the application moved to every rank (`id128`), and rgbdfe_verify_sharded_enqueue / _collect."""

    def __init__(self, device, rank, world, n_kf_total, kf_desc_block, kf_xyz_block, id128=None, max_feats=1000, depth_cov=(0.01 * 2.0 * 2.0) ** 2):
        import rgbdfe

        self.rgbdfe, self.rank, self.world, self.n_kf = rgbdfe, rank, world, n_kf_total
        self.lo, self.hi = rgbdfe.shard_range(n_kf_total, rank, world)
        assert len(kf_desc_block) == self.hi - self.lo
        self.ctx = rgbdfe.Context(device, 640, 480, max_feats, 1)
        self.ctx.depth_cov_set(1, depth_cov)  # the latch must be pinned: results then do not depend on pair order
        self.db = self.ctx.kfdb_create(kf_desc_block, kf_xyz_block)
        self.comm = self.ctx.nccl_comm_create(world, rank, id128) if world > 1 else None
        self.params = rgbdfe.VerifyParams(0.9, 30, 200, 30, 3.0, 4, 42)

    def enqueue(self, q_desc, q_xyz, nq, n_queries, slot, seed=42):
        self.params.seed = seed
        self.ctx.verify_sharded_enqueue(self.db, self.comm, self.world, self.rank, self.n_kf, q_desc, q_xyz, nq, n_queries, self.params, slot)

    def collect(self, slot, n_queries, want=True, out=None):
        return self.ctx.verify_sharded_collect(slot, n_queries, self.n_kf, want, out)

    def verify(self, q_desc, q_xyz, seed=42):
        q_desc = np.ascontiguousarray(q_desc, np.uint8)
        q_xyz = np.ascontiguousarray(q_xyz, np.float32)
        nq = q_desc.shape[-2]
        n_queries = 1 if q_desc.ndim == 2 else q_desc.shape[0]
        self.enqueue(q_desc, q_xyz, nq, n_queries, 0, seed)
        return self.collect(0, n_queries)

    def close(self):
        if self.comm:
            self.rgbdfe.lib().rgbdfe_nccl_comm_destroy(self.comm)
            self.comm = None
        if self.ctx:
            self.ctx.kfdb_destroy(self.db)
            self.ctx.close()
            self.ctx = None
