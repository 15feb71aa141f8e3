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


def gather_results(local, n_kf, rank, world, device="cpu"):
    """all_gather the per-pair result structs of every rank into keyframe order.  `local`: structured array of
    this rank's block (or a device uint8 tensor with the same bytes).  Returns a uint8 torch tensor [n_kf, 84]."""
    import torch
    import torch.distributed as dist

    per = -(-n_kf // world)  # equal-size slots for all_gather_into_tensor; short blocks are padded
    if isinstance(local, np.ndarray):
        buf = torch.zeros(per * RESULT_BYTES, dtype=torch.uint8, device=device)
        raw = torch.from_numpy(np.ascontiguousarray(local).view(np.uint8).reshape(-1).copy())
        buf[: raw.numel()] = raw.to(device)
    else:
        buf = local
        assert buf.numel() == per * RESULT_BYTES
    if world == 1:
        return buf.view(per, RESULT_BYTES)[:n_kf]
    out = torch.empty(world * per * RESULT_BYTES, dtype=torch.uint8, device=buf.device)
    dist.all_gather_into_tensor(out, buf)
    out = out.view(world, per, RESULT_BYTES)
    rows = []
    for r in range(world):
        lo, hi = shard_range(n_kf, r, world)
        rows.append(out[r, : hi - lo])
    return torch.cat(rows, 0)


def first_success(results, min_index=0):
    """detectLoop takes the FIRST candidate, in candidate order, whose RansacSE3 succeeds (PoseGraph.cpp:283)."""
    ok = np.nonzero(results["ok"][min_index:] != 0)[0]
    return int(ok[0]) + min_index if len(ok) else -1


class ShardedVerifier:
    """Keyframe block of this rank + the verify call.  `local_verify(kf_lo, kf_hi, q_desc, q_xyz, seed)` must return
    the structured results of keyframes [kf_lo, kf_hi) with per-pair stream srand(seed + global index)."""

    def __init__(self, n_kf, rank, world, local_verify, device="cpu"):
        self.n_kf, self.rank, self.world, self.device = n_kf, rank, world, device
        self.lo, self.hi = shard_range(n_kf, rank, world)
        self.local_verify = local_verify

    def verify_many(self, q_descs, q_xyzs, seed, dtype):
        """Batched relocalisation (BASELINE configs[4]: Q query frames x n_kf keyframes, all pairs): every rank verifies
        all Q queries against its keyframe block, then ONE all_gather moves the Q x block results.  Pair (q, k) uses the
        stream srand(seed + q * n_kf + k), so the result is independent of the sharding.  Returns [Q, n_kf] records."""
        import torch

        Q = len(q_descs)
        per = -(-self.n_kf // self.world)
        buf = torch.zeros(Q * per * RESULT_BYTES, dtype=torch.uint8)
        view = buf.view(Q, per * RESULT_BYTES)
        for q in range(Q):
            local = self.local_verify(self.lo, self.hi, q_descs[q], q_xyzs[q], seed + q * self.n_kf)
            raw = torch.from_numpy(np.ascontiguousarray(local).view(np.uint8).reshape(-1).copy())
            view[q, : raw.numel()] = raw
        buf = buf.to(self.device)
        if self.world == 1:
            out = buf.view(1, Q, per, RESULT_BYTES)
        else:
            import torch.distributed as dist

            out = torch.empty(self.world * buf.numel(), dtype=torch.uint8, device=buf.device)
            dist.all_gather_into_tensor(out, buf)
            out = out.view(self.world, Q, per, RESULT_BYTES)
        rows = []
        for r in range(self.world):
            lo, hi = shard_range(self.n_kf, r, self.world)
            rows.append(out[r, :, : hi - lo])
        g = torch.cat(rows, 1).contiguous()  # [Q, n_kf, 84]
        return np.frombuffer(g.cpu().numpy().tobytes(), dtype=dtype).reshape(Q, self.n_kf)

    def verify(self, q_desc, q_xyz, seed, dtype):
        local = self.local_verify(self.lo, self.hi, q_desc, q_xyz, seed)
        g = gather_results(local, self.n_kf, self.rank, self.world, self.device)
        return np.frombuffer(g.cpu().numpy().tobytes(), dtype=dtype)
