// K9-K11 — RansacSE3::compute (Solver/SolverSE3.cpp:23-133) for a batch of pairs.
//
// The reference loop is sequential only in (a) the rand() stream and (b) the accept /
// "n += 10" / break control; each iteration's refinement result depends on nothing but
// its sample.  So:
//   k_se3_prepare    1 CTA / pair: sorts the matches into exactly the permutation libstdc++'s
//                    std::sort produces (std_sort.cuh), gathers the matched 3D points and TFC
//                    weights in sorted order, latches the depth covariance, marks F2 outliers,
//                    and (one warp, concurrently with the sort) draws the biased 4-samples of
//                    the first chunk of iterations from the glibc-compatible stream — samples
//                    depend only on the match count, never on the data.
//   k_se3_hypotheses 1 CTA / (pair, iteration): up to 19 refinement rounds of
//                    {TFC fit (FP32) -> Mahalanobis scoring of all matches (FP64, all threads)
//                    -> ordered inlier list -> ordered FP64 error sum}.  A finished hypothesis
//                    publishes its result and then replays SolverSE3.cpp:86-101 in iteration
//                    order over everything published so far (one CTA at a time, under a
//                    per-pair lock).  Once the replay sees the loop end (break / iterations
//                    used up) the pair is finalised right there — best model, identity
//                    fallback (:105-117), Frame flags (:119-123), rand() stream advanced by
//                    exactly the calls of the executed iterations — and hypotheses still
//                    running for that pair stop at their next round: the chunk costs the
//                    latency of the iterations the reference would have run, not of its
//                    slowest member.  If the chunk is used up the same CTA draws the next
//                    chunk's samples.
// Hypotheses are issued in growing chunks; chunks of a finished pair exit immediately.
#include "match_filter.cuh"
#include "rng.cuh"
#include "se3_math.cuh"
#include "std_sort.cuh"

namespace rgbdfe {

#define SE3_MAX_SAMPLE 8
#define PREP_THREADS 256
#define RNG_BLOCK 512 // rand() outputs generated per refill of the sampler's window

// smem scratch of the sampler / stream roll: unsigned lin[34 + RNG_BLOCK]; int ids[RNG_BLOCK]
#define RNG_SCRATCH_WORDS (RNG_HIST + 2 * RNG_BLOCK)

__device__ __forceinline__ int ld_volatile(const int* p) { return *reinterpret_cast<const volatile int*>(p); }

// sampleMatches (SolverSE3.cpp:135-159) for iterations [k0, k1) of one pair; one warp.
// Lane 0 extends the rand() window and walks the set-insert logic; all lanes convert the raw
// words to Random::randomInt(0, M-1) values.  `st` is read and left at the position after the
// last consumed rand() call.
__device__ void warp_sample_chunk(unsigned* lin, int* ids, RngState* st, int M, int ss, int k0, int k1, int* __restrict__ samples,
    int* __restrict__ used, int& calls, int lane)
{
    int idx0 = 0;
    if (lane == 0)
        idx0 = rng_window_load(st, lin);
    int k = k0, cnt = 0, safetyNet = 0;
    int cur[SE3_MAX_SAMPLE];
    int total = 0; // outputs consumed since *st was read
    while (k < k1) {
        // expected 2 * ss draws per iteration; 25 % head room, whole pairs
        int nb = 2 * ss * (k1 - k);
        nb = min(RNG_BLOCK, (nb + (nb >> 2) + 8) & ~1);
        if (lane == 0)
            rng_extend(lin, nb);
        __syncwarp();
        for (int t = lane; t < nb; t += 32)
            ids[t] = rng_random_int0(lin[RNG_HIST + t], M);
        __syncwarp();
        int pos = 0;
        if (lane == 0) {
            while (k < k1 && pos + 2 <= nb) {
                int id1 = ids[pos];
                const int id2 = ids[pos + 1];
                pos += 2;
                calls += 2;
                if (id1 > id2)
                    id1 = id2;
                int p = 0; // std::set insert: ascending, duplicates ignored
                while (p < cnt && cur[p] < id1)
                    p++;
                if (!(p < cnt && cur[p] == id1)) {
                    for (int j = cnt; j > p; j--)
                        cur[j] = cur[j - 1];
                    cur[p] = id1;
                    cnt++;
                }
                ++safetyNet;
                if (cnt >= ss || safetyNet > 10000) {
                    for (int j = 0; j < SE3_MAX_SAMPLE; j++)
                        samples[k * SE3_MAX_SAMPLE + j] = j < cnt ? cur[j] : -1;
                    used[k] = calls;
                    k++;
                    cnt = 0;
                    safetyNet = 0;
                }
            }
            // slide the window: the state after `pos` outputs is lin[pos .. pos+34)
            for (int t = 0; t < RNG_HIST; t++)
                lin[t] = lin[pos + t];
        }
        k = __shfl_sync(0xffffffffu, k, 0);
        pos = __shfl_sync(0xffffffffu, pos, 0);
        total += pos;
        __syncwarp();
    }
    if (lane == 0)
        rng_window_store(st, lin, idx0 + total);
    calls = __shfl_sync(0xffffffffu, calls, 0);
    __syncwarp();
}

// dynamic smem: SKey keys[cap]; int Lpos[cap], Rpos[cap], leaves[cap]; int3 qa[QCAP], qb[QCAP]
__global__ void __launch_bounds__(PREP_THREADS) k_se3_prepare(const MatchPair* __restrict__ pairs, int slot0, FusedFilter ff,
    int4* partial, rgbdfe_dmatch* matches_all,
    rgbdfe_dmatch* __restrict__ sorted_all, rgbdfe_point3f* __restrict__ P1_all, rgbdfe_point3f* __restrict__ P2_all, float* __restrict__ w_all,
    int match_cap, int* __restrict__ nmatch, Se3Params prm, int update_f2, int per_pair_rng, unsigned seed, int first_pair_index,
    int min_matches, const RngState* __restrict__ ctx_rng, RngState* __restrict__ pair_rng, DepthLatch* __restrict__ latch,
    int* __restrict__ samples_all, int* __restrict__ rng_used_all, ReplayState* __restrict__ replay, int* __restrict__ ready_all,
    int* __restrict__ lock_all, int first_chunk, long long* dbg)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SKey* keys = reinterpret_cast<SKey*>(smem_raw);
    int* Lpos = reinterpret_cast<int*>(keys + match_cap);
    int* Rpos = Lpos + match_cap;
    int* leaves = Rpos + match_cap;
    int3* qa = reinterpret_cast<int3*>(leaves + match_cap);
    int3* qb = qa + SORT_QCAP;
    __shared__ int s_counters[4];
    __shared__ unsigned s_rng[RNG_SCRATCH_WORDS];

    const int pair = slot0 + blockIdx.x;
    const MatchPair P = pairs[pair];
    const rgbdfe_dmatch* m = matches_all + (size_t)pair * match_cap;
    rgbdfe_dmatch* sorted = sorted_all + (size_t)pair * match_cap;
    const int tid = threadIdx.x;
    if (blockIdx.x != 0 || tid != 0)
        dbg = nullptr;
    DBG_STAMP(dbg, 0);
    int M;
    if (ff.splits > 0) {
        // Matcher::match's filter half (K8) on this pair's kNN partials, fused into the chain
        M = cta_ratio_filter(P, pair, reinterpret_cast<int*>(smem_raw), partial, match_cap, ff.splits, ff.ratio, ff.discard_outliers,
            matches_all, match_cap, nullptr);
        if (tid == 0)
            nmatch[pair] = M;
        __syncthreads();
    } else
        M = nmatch[pair];
    DBG_STAMP(dbg, 1);
    const bool run = M >= (int)prm.min_inlier_th && M >= min_matches; // SolverSE3.cpp:29 / PoseGraph.cpp:142-144
    const int ss = (int)prm.sample_size;
    const int n_first = (M >= ss) ? min(min(prm.iterations, MAX_ITERS), first_chunk) : 0;
    for (int i = tid; i < MAX_ITERS; i += PREP_THREADS)
        ready_all[(size_t)pair * MAX_ITERS + i] = 0;
    if (tid == 0) {
        lock_all[pair] = 0;
        ReplayState st;
        st.n = 0;
        st.k = 0;
        st.done = 0;
        st.validIters = 0;
        st.bestHyp = -1;
        st.bestCount = 0;
        st.rmse = 1e6f;
        st.run = run ? 1 : 0;
        st.nm = M;
        st.rng_calls = 0;
        st.sampled_k = run ? n_first : 0;
        st.sampled_calls = 0; // filled by the sampler warp below
        replay[pair] = st;
    }
    if (!run)
        return;
    for (int i = tid; i < M; i += PREP_THREADS) {
        const rgbdfe_dmatch mm = m[i];
        keys[i].d = mm.distance;
        keys[i].i = i;
        // SolverSE3.cpp:38-42: every matched train keypoint of F2 becomes an outlier
        if (update_f2 && P.t_outlier)
            P.t_outlier[mm.trainIdx] = 1;
    }
    __syncthreads();
    // the sampler needs only M: the last warp draws the first chunk while the others sort
    if (tid >= PREP_THREADS - 32) {
        const int lane = tid & 31;
        RngState* s = pair_rng + pair; // working copy, continued by the replay when more chunks are needed
        if (lane == 0) {
            if (per_pair_rng)
                rng_seed_state(*s, seed + (unsigned)(first_pair_index + (ff.seed_group > 0 ? ((int)blockIdx.x / ff.seed_group) * ff.seed_stride + (int)blockIdx.x % ff.seed_group : (int)blockIdx.x)), s_rng);
            else
                *s = *ctx_rng;
        }
        __syncwarp();
        int calls = 0;
        warp_sample_chunk(s_rng, reinterpret_cast<int*>(s_rng + RNG_HIST + RNG_BLOCK), s, M, ss, 0, n_first,
            samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE, rng_used_all + (size_t)pair * MAX_ITERS, calls, lane);
        if (lane == 0)
            replay[pair].sampled_calls = calls;
    } else
        cta_std_sort(keys, M, Lpos, Rpos, leaves, qa, qb, s_counters, PREP_THREADS - 32, dbg ? dbg + 400 : nullptr); // :52
    DBG_STAMP(dbg, 2);
    __syncthreads();
    DBG_STAMP(dbg, 3);
    rgbdfe_point3f* P1 = P1_all + (size_t)pair * match_cap;
    rgbdfe_point3f* P2 = P2_all + (size_t)pair * match_cap;
    float* w = w_all + (size_t)pair * match_cap;
    for (int j = tid; j < M; j += PREP_THREADS) {
        const rgbdfe_dmatch mm = m[keys[j].i];
        sorted[j] = mm;
        const rgbdfe_point3f from = P.q_xyz[mm.queryIdx], to = P.t_xyz[mm.trainIdx];
        P1[j] = from;
        P2[j] = to;
        // TFC weight (SolverSE3.cpp:172-176); 0 encodes "skipped" (NaN depth, or tfc.add's weight == 0 early return)
        float wt = 1.0f / (from.z * to.z);
        if (isnan(from.z) || isnan(to.z) || wt == 0.0f)
            wt = 0.0f; // canonical +0
        w[j] = wt;
    }
    __syncthreads();
    DBG_STAMP(dbg, 4);
    // depthCovariance latch (:282-287): the first source depth that reaches errorFunction2
    if (tid < 32 && !latch->latched) {
        for (int j0 = 0; j0 < M; j0 += 32) {
            const int j = j0 + tid;
            bool hit = false;
            rgbdfe_point3f o = { 0.f, 0.f, 0.f };
            if (j < M) {
                o = P1[j];
                const rgbdfe_point3f t = P2[j];
                hit = !(o.z == 0.0f || t.x == 0.0f) && !(isnan(o.z) || isnan(t.z));
            }
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (bal) {
                if (tid == __ffs(bal) - 1) {
                    const double z = (double)o.z;
                    const double sd = 0.01 * z * z;
                    latch->cov = sd * sd;
                    __threadfence();
                    latch->latched = 1;
                }
                break;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// Sequential control of the RANSAC loop, replayed over published hypotheses (one warp,
// uniform control flow; lane 0 owns the stores).
// ---------------------------------------------------------------------------------------
struct ReplayCtx {
    const MatchPair* pairs;
    const rgbdfe_dmatch* sorted_all;
    const rgbdfe_point3f *P1_all, *P2_all;
    int match_cap;
    Se3Params prm;
    int update_f2, per_pair_rng;
    RngState *ctx_rng, *pair_rng;
    const DepthLatch* latch;
    int *samples_all, *rng_used_all;
    const HypOut* hyp_all;
    const unsigned* mask_all;
    int mask_words;
    ReplayState* replay;
    int *ready_all, *lock_all;
    rgbdfe_se3_result* results; // results[pair - slot0]
    int slot0;
    int* inl_pos_all;
    long long* dbg;
};

__device__ __forceinline__ ReplayState load_state(const ReplayState* p)
{
    const volatile int* v = reinterpret_cast<const volatile int*>(p);
    ReplayState st;
    int* d = reinterpret_cast<int*>(&st);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(ReplayState) / sizeof(int)); i++)
        d[i] = v[i];
    return st;
}
__device__ __forceinline__ void store_state(ReplayState* p, const ReplayState& st)
{
    volatile int* v = reinterpret_cast<volatile int*>(p);
    const int* s = reinterpret_cast<const int*>(&st);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(ReplayState) / sizeof(int)); i++)
        v[i] = s[i];
}

// Loop ended: pick the model, identity fallback, Frame flags, result, stream position.
// scratch: unsigned[RNG_SCRATCH_WORDS + mask_words] of shared memory.
__device__ void warp_finalize(const ReplayCtx& R, int pair, ReplayState& st, unsigned* scratch, int lane)
{
    const Se3Params& prm = R.prm;
    const int M = st.nm, match_cap = R.match_cap;
    const MatchPair P = R.pairs[pair];
    const rgbdfe_dmatch* m = R.sorted_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P1 = R.P1_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P2 = R.P2_all + (size_t)pair * match_cap;
    rgbdfe_se3_result* res = R.results + (pair - R.slot0);
    int* inl_pos = R.inl_pos_all + (size_t)pair * match_cap;
    unsigned* s_mask = scratch + RNG_SCRATCH_WORDS;

    st.rng_calls = st.k > 0 ? R.rng_used_all[(size_t)pair * MAX_ITERS + st.k - 1] : 0;
    if (!R.per_pair_rng && lane == 0)
        rng_advance(scratch, R.ctx_rng, st.rng_calls, 2 * RNG_BLOCK); // the process stream moves by what the executed iterations drew
    float T[16];
#pragma unroll
    for (int i = 0; i < 16; i++)
        T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    int n_inl = 0;
    float rmse = st.rmse;
    const unsigned* mask = nullptr;
    bool mask_global = false;
    const int words = (M + 31) >> 5;
    if (st.validIters == 0) {
        // identity fallback (SolverSE3.cpp:105-117)
        const double th = (double)(prm.max_mahal * prm.max_mahal);
        double sum = 0.0;
        int cnt = 0;
        double Td[16];
#pragma unroll
        for (int i = 0; i < 16; i++)
            Td[i] = (double)T[i];
        const double C = R.latch->cov;
        for (int w = 0; w < words; w++) {
            const int j = w * 32 + lane;
            double d2 = 0.0;
            bool in = false;
            if (j < M) {
                const rgbdfe_point3f o = P1[j], t = P2[j];
                if (!(o.z == 0.0f || t.x == 0.0f)) {
                    const double md = mahal2(o, t, Td, C, prm.rcx, prm.rcy);
                    if (!(md > th) && (md >= 0.0)) {
                        in = true;
                        d2 = md;
                    }
                }
            }
            unsigned bal = __ballot_sync(0xffffffffu, in);
            if (lane == 0)
                s_mask[w] = bal;
            cnt += __popc(bal);
            while (bal) {
                const int l = __ffs(bal) - 1;
                bal &= bal - 1;
                sum += __shfl_sync(0xffffffffu, d2, l);
            }
        }
        __syncwarp();
        double err = 1e9;
        if (cnt >= 3)
            err = sqrt(sum / (double)cnt);
        if (cnt > (int)prm.min_inlier_th && err < (double)prm.max_mahal) {
            n_inl = cnt;
            rmse = (float)((double)rmse + err);
            mask = s_mask;
        }
    } else if (st.bestHyp >= 0) {
        const HypOut* h = R.hyp_all + (size_t)pair * MAX_ITERS + st.bestHyp;
#pragma unroll
        for (int i = 0; i < 16; i++)
            T[i] = __ldcg(&h->T[i]);
        n_inl = st.bestCount;
        mask = R.mask_all + ((size_t)pair * MAX_ITERS + st.bestHyp) * R.mask_words;
        mask_global = true;
    }
    const int ok = n_inl >= (int)prm.min_inlier_th ? 1 : 0;
    // inlier positions (sorted-match order) + Frame::setInlier (:119-123)
    if (mask && n_inl > 0) {
        int base = 0;
        for (int w = 0; w < words; w++) {
            const unsigned bits = mask_global ? __ldcg(&mask[w]) : mask[w];
            if (bits & (1u << lane)) {
                const int j = w * 32 + lane;
                inl_pos[base + __popc(bits & ((1u << lane) - 1))] = j;
                if (ok && R.update_f2 && P.t_outlier)
                    P.t_outlier[m[j].trainIdx] = 0;
            }
            base += __popc(bits);
        }
    }
    if (lane == 0) {
        res->ok = ok;
        res->rmse = rmse;
        for (int i = 0; i < 16; i++)
            res->T21[i] = T[i];
        res->n_inliers = n_inl;
        res->iters_executed = st.k;
        res->n_matches = M;
    }
}

// Called by warp 0 of a CTA (all 32 lanes).  chunk_end: hypotheses [.., chunk_end) belong to this
// launch; next_end: where the next chunk will end (samples are drawn up to there when the loop
// goes on).
__device__ void warp_try_replay(const ReplayCtx& R, int pair, int chunk_end, int next_end, unsigned* scratch, int lane)
{
    ReplayState* rp = R.replay + pair;
    const int* ready = R.ready_all + (size_t)pair * MAX_ITERS;
    int* lock = R.lock_all + pair;
    const Se3Params& prm = R.prm;
    for (;;) {
        int got = 0;
        if (lane == 0)
            got = atomicCAS(lock, 0, 1) == 0;
        got = __shfl_sync(0xffffffffu, got, 0);
        if (!got)
            return; // the holder re-checks the ready flags after it releases
        __threadfence();
        ReplayState st = load_state(rp);
        if (st.done == 2) {
            if (lane == 0)
                atomicExch(lock, 0);
            return;
        }
        const int M = st.nm;
        if (!st.run) {
            if (lane == 0) {
                rgbdfe_se3_result* res = R.results + (pair - R.slot0);
                res->ok = 0;
                res->rmse = 1e6f;
                for (int i = 0; i < 16; i++)
                    res->T21[i] = (i % 5 == 0) ? 1.0f : 0.0f;
                res->n_inliers = 0;
                res->iters_executed = 0;
                res->n_matches = M;
                st.done = 2;
                store_state(rp, st);
                __threadfence();
                atomicExch(lock, 0);
            }
            return;
        }
        const bool Mok = M >= (int)prm.sample_size;
        const int max_iters = min(prm.iterations, MAX_ITERS);
        if (!st.done) {
            bool broke = false;
            const HypOut* hyp = R.hyp_all + (size_t)pair * MAX_ITERS;
            while (st.n < max_iters && Mok && st.k < chunk_end && st.k < st.sampled_k) {
                if (!ld_volatile(&ready[st.k]))
                    break;
                __threadfence();
                const int cnt = __ldcg(&hyp[st.k].count);
                const double err = __ldcg(&hyp[st.k].err);
                const int k = st.k;
                st.k++;
                if (cnt > 0) {
                    st.validIters++;
                    if (err <= (double)st.rmse && cnt >= st.bestCount && cnt >= (int)prm.min_inlier_th) {
                        st.rmse = (float)err;
                        st.bestHyp = k;
                        st.bestCount = cnt;
                        if ((double)cnt > M * 0.5)
                            st.n += 10;
                        if ((double)cnt > M * 0.75)
                            st.n += 10;
                        if ((double)cnt > M * 0.8) {
                            broke = true;
                            break;
                        }
                    }
                }
                st.n++;
            }
            if (broke || st.n >= max_iters || !Mok)
                st.done = 1;
        }
        if (st.done) {
            if (lane == 0)
                *reinterpret_cast<volatile int*>(&rp->done) = 1; // running hypotheses of this pair stop at their next round
            warp_finalize(R, pair, st, scratch, lane);
            st.done = 2;
            if (lane == 0) {
                store_state(rp, st);
                __threadfence();
                atomicExch(lock, 0);
            }
            return;
        }
        // the loop goes on; once this chunk is used up, draw the next chunk's samples (continuing the pair's stream)
        const int k1 = min(next_end >= 0 ? next_end : st.sampled_k - next_end, max_iters); // next_end < 0: draw -next_end more iterations
        if (st.k >= st.sampled_k && k1 > st.sampled_k) {
            int calls = st.sampled_calls;
            warp_sample_chunk(scratch, reinterpret_cast<int*>(scratch + RNG_HIST + RNG_BLOCK), R.pair_rng + pair, M, (int)prm.sample_size,
                st.sampled_k, k1, R.samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE, R.rng_used_all + (size_t)pair * MAX_ITERS, calls, lane);
            st.sampled_calls = calls;
            st.sampled_k = k1;
        }
        if (lane == 0) {
            store_state(rp, st);
            __threadfence();
            atomicExch(lock, 0);
            __threadfence();
        }
        __syncwarp();
        // a hypothesis may have been published while the lock was held (its CTA then gave up): look again
        const bool again = st.k < chunk_end && st.k < st.sampled_k && ld_volatile(&ready[st.k]);
        if (!again)
            return;
    }
}

// Draw the samples of iterations [sampled_k, sampled_k + chunk) of a pair ahead of their consumption, so that iteration
// `need` exists.  Called by one warp that would otherwise wait for the replay to draw them; takes the pair lock (held
// only by resident, running warps) or backs off for a moment.  scratch: as for warp_try_replay.
__device__ void warp_draw_ahead(const ReplayCtx& R, int pair, int need, int chunk, unsigned* scratch, int lane)
{
    ReplayState* rp = R.replay + pair;
    int* lock = R.lock_all + pair;
    int got = 0;
    if (lane == 0)
        got = atomicCAS(lock, 0, 1) == 0;
    got = __shfl_sync(0xffffffffu, got, 0);
    if (!got) {
        __nanosleep(200);
        return;
    }
    __threadfence();
    ReplayState st = load_state(rp);
    const int max_iters = min(R.prm.iterations, MAX_ITERS);
    if (!st.done && st.run && st.sampled_k <= need && st.sampled_k < max_iters && st.nm >= (int)R.prm.sample_size) {
        const int k1 = min(max(need + 1, st.sampled_k + chunk), max_iters);
        int calls = st.sampled_calls;
        warp_sample_chunk(scratch, reinterpret_cast<int*>(scratch + RNG_HIST + RNG_BLOCK), R.pair_rng + pair, st.nm, (int)R.prm.sample_size, st.sampled_k,
            k1, R.samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE, R.rng_used_all + (size_t)pair * MAX_ITERS, calls, lane);
        st.sampled_calls = calls;
        st.sampled_k = k1;
        if (lane == 0) {
            // only the two sampler fields changed; the replay fields are re-read under the lock by whoever holds it next
            *reinterpret_cast<volatile int*>(&rp->sampled_calls) = st.sampled_calls;
            __threadfence();
            *reinterpret_cast<volatile int*>(&rp->sampled_k) = st.sampled_k;
        }
    }
    if (lane == 0) {
        __threadfence();
        atomicExch(lock, 0);
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------
// One hypothesis = one CTA.  Shared memory (dynamic), capA = cap rounded up to 4:
//   double g_d2[capA]                 d^2 of the current inliers, list order
//   float  g_w[capA], g_alpha[capA]   TFC weight / accumulated weight -> alpha, list order
//   float  g_f[3][capA], g_t[3][capA] matched points of the current inliers (SoA), list order
//   unsigned maskA[mw], maskB[mw]     inlier bitmasks (refined / current)
//   int wbase[mw]                     per-word inlier counts -> exclusive prefix
// The same bytes are reused as scratch by the replay (RNG window, identity-fallback mask).
// ---------------------------------------------------------------------------------------
template <int THREADS>
__device__ __forceinline__ void group_bar()
{
    asm volatile("bar.sync 1, %0;" ::"r"(THREADS - 32) : "memory");
}

// pcl::TransformationFromCorrespondences over the gathered list (SolverSE3.cpp:161-179), run by
// every warp except warp 1.  The reference's `add` is one sequential recurrence per point; here
//   (1) thread 0 forms the running weight sums in list order (one dependent FADD per point),
//   (2) all threads turn them into alpha = w / accW (the division leaves the critical chain),
//   (3) warp 0 runs the mean / covariance recurrences: lane c < 9 owns cov(c/3, c%3) and tracks
//       mean2[c/3], mean1[c%3] itself, so every value goes through exactly the operations of the
//       sequential `add` while the dependent chain per point is sub-mul-add.
// Points the reference skips (NaN depth, weight == 0) are stored as w = +0, alpha = 0, x = y = 0:
// accW + 0, mean + 0 * d and 1 * (cov + 0) leave every accumulator unchanged, so the loops are
// branch-free.  The list is zero-padded to a multiple of 4 (+4 for the prefetch).
// ctl: when non-null, the sequential loops poll it and give up as soon as the round's decision is "stop".
template <int THREADS>
__device__ void group_tfc_fit(const float* g_w, float* g_alpha, const float* g_f, const float* g_t, int capA, int cnt, float* Tout,
    const volatile int* ctl, int tid, long long* dbg = nullptr)
{
    const int lane = tid & 31, warp = tid >> 5;
    if (warp == 0) {
        // all 32 lanes run the same recurrence (same addresses, same values): a warp without a divergent
        // sibling path issues this loop back to back, one lane alone does not
        float acc = 0.0f;
        float4 w4 = *reinterpret_cast<const float4*>(g_w);
        for (int k = 0; k < cnt; k += 4) {
            const float4 wn = *reinterpret_cast<const float4*>(g_w + k + 4);
            float4 a4;
            acc += w4.x;
            a4.x = acc;
            acc += w4.y;
            a4.y = acc;
            acc += w4.z;
            a4.z = acc;
            acc += w4.w;
            a4.w = acc;
            *reinterpret_cast<float4*>(g_alpha + k) = a4;
            w4 = wn;
        }
    }
    DBG_STAMP(dbg, 0);
    group_bar<THREADS>();
    const int gtid = warp == 0 ? lane : tid - 32;
    for (int k = gtid; k < cnt; k += THREADS - 32) {
        const float w = g_w[k];
        g_alpha[k] = (w != 0.0f) ? w / g_alpha[k] : 0.0f;
    }
    for (int k = cnt + gtid; k < ((cnt + 3) & ~3) + 4; k += THREADS - 32)
        g_alpha[k] = 0.0f;
    group_bar<THREADS>();
    DBG_STAMP(dbg, 1);
    if (warp != 0)
        return;
    const int c = lane % 9, ci = c / 3, cj = c % 3;
    const float* gf = g_f + cj * capA;
    const float* gt = g_t + ci * capA;
    float m1 = 0.0f, m2 = 0.0f, cv = 0.0f;
    for (int k0 = 0; k0 < cnt; k0 += 128) {
        if (ctl) { // polled between blocks only, so that the inner loop stays a plain counted loop
            const int v = *ctl;
            if (v >= 0 && (v & 1))
                return; // this round stops the refinement: the fit would be discarded
        }
        const int k1 = min(cnt, k0 + 128);
        float4 a4 = *reinterpret_cast<const float4*>(g_alpha + k0);
        float4 x4 = *reinterpret_cast<const float4*>(gf + k0);
        float4 y4 = *reinterpret_cast<const float4*>(gt + k0);
        for (int k = k0; k < k1; k += 4) {
            const float4 an = *reinterpret_cast<const float4*>(g_alpha + k + 4);
            const float4 xn = *reinterpret_cast<const float4*>(gf + k + 4);
            const float4 yn = *reinterpret_cast<const float4*>(gt + k + 4);
            const float aa[4] = { a4.x, a4.y, a4.z, a4.w };
            const float xx[4] = { x4.x, x4.y, x4.z, x4.w }, yy[4] = { y4.x, y4.y, y4.z, y4.w };
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float alpha = aa[u];
                const float d1 = xx[u] - m1, d2 = yy[u] - m2;
                const float oma = 1.0f - alpha;
                const float ad2 = alpha * d2;
                cv = oma * (cv + ad2 * d1);
                m1 += alpha * d1;
                m2 += ad2;
            }
            a4 = an;
            x4 = xn;
            y4 = yn;
        }
    }
    DBG_STAMP(dbg, 2);
    Tfc tfc;
    tfc.accW = 0.0f;
#pragma unroll
    for (int e = 0; e < 9; e++)
        tfc.cov[e] = __shfl_sync(0xffffffffu, cv, e);
#pragma unroll
    for (int e = 0; e < 3; e++) {
        tfc.mean1[e] = __shfl_sync(0xffffffffu, m1, e);     // lanes (0, j)
        tfc.mean2[e] = __shfl_sync(0xffffffffu, m2, 3 * e); // lanes (i, 0)
    }
    float T[16];
    tfc.transform(T);
    if (lane < 16)
        Tout[lane] = T[lane];
}

// One hypothesis (iteration h of one pair) evaluated by the whole CTA; `lead` marks the CTA that drives the replay when
// there is nothing to evaluate.
template <int THREADS, int NPT>
__device__ __forceinline__ void cta_se3_hypothesis(const ReplayCtx& R, const float* __restrict__ w_all, int pair, int h, int hyp_begin, int hyp_end,
    int next_end, bool lead)
{
    constexpr int NW = THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int match_cap = R.match_cap, capA = ((match_cap + 3) & ~3) + 8, mw = R.mask_words;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* g_d2 = reinterpret_cast<double*>(smem_raw);
    float* g_w = reinterpret_cast<float*>(g_d2 + capA);
    float* g_alpha = g_w + capA;
    float* g_f = g_alpha + capA;
    float* g_t = g_f + 3 * capA;
    unsigned* maskA = reinterpret_cast<unsigned*>(g_t + 3 * capA);
    unsigned* maskB = maskA + mw;
    int* wbase = reinterpret_cast<int*>(maskB + mw);
    unsigned* scratch = reinterpret_cast<unsigned*>(smem_raw);
    __shared__ float sT[2][16], sRefT[16];
    __shared__ int s_active, s_M, s_cnt, s_ctl, s_abort, s_refCount;
    __shared__ double s_refErr;

    if (tid == 0) {
        const ReplayState* rp = R.replay + pair;
        const int done = ld_volatile(&rp->done), run = ld_volatile(&rp->run), sk = ld_volatile(&rp->sampled_k);
        s_active = (!done && run && h < hyp_end && h < sk) ? 1 : 0;
        s_M = ld_volatile(&rp->nm);
        s_abort = 0;
        s_refCount = 0;
        s_refErr = 1e6;
    }
    __syncthreads();
    if (!s_active) {
        // nothing to evaluate; the first CTA of the launch still drives the replay (pairs without a loop, exhausted iterations)
        if (lead && warp == 0)
            warp_try_replay(R, pair, hyp_end, next_end, scratch, lane);
        return;
    }
    const int M = s_M, words = (M + 31) >> 5;
    long long* dbg = (h == 0 && blockIdx.y == 0 && tid == 0) ? R.dbg : nullptr;
    DBG_STAMP(dbg, 16);
    if (dbg)
        dbg[300] = clock64();
    const rgbdfe_point3f* P1 = R.P1_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P2 = R.P2_all + (size_t)pair * match_cap;
    const float* P1f = reinterpret_cast<const float*>(P1);
    const float* P2f = reinterpret_cast<const float*>(P2);
    const float* w = w_all + (size_t)pair * match_cap;
    const Se3Params prm = R.prm;
    const double C = R.latch->cov;
    const double th = (double)(prm.max_mahal * prm.max_mahal);

    auto gather = [&](int k, int j) {
        const float wj = w[j]; // +0 marks a correspondence the fit skips
        g_w[k] = wj;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            g_f[c * capA + k] = wj != 0.0f ? P1f[j * 3 + c] : 0.0f;
            g_t[c * capA + k] = wj != 0.0f ? P2f[j * 3 + c] : 0.0f;
        }
    };
    auto pad_list = [&](int cnt) { // zero entries [cnt, roundup4(cnt) + 4): the fit loops run in fours and prefetch one step
        const int end = ((cnt + 3) & ~3) + 4;
        for (int k = cnt + tid; k < end; k += THREADS) {
            g_w[k] = 0.0f;
#pragma unroll
            for (int c = 0; c < 3; c++)
                g_f[c * capA + k] = g_t[c * capA + k] = 0.0f;
        }
    };

    unsigned* ref_m = maskA;
    unsigned* cur_m = maskB;
    for (int i = tid; i < words; i += THREADS)
        maskA[i] = 0;
    if (tid < 16)
        sRefT[tid] = (tid % 5 == 0) ? 1.0f : 0.0f;
    {
        const int* s = R.samples_all + ((size_t)pair * MAX_ITERS + h) * SE3_MAX_SAMPLE;
        if (tid < SE3_MAX_SAMPLE) {
            const int j = s[tid]; // ascending ids first, then -1 padding
            if (j >= 0)
                gather(tid, j);
        }
        if (tid == 0) {
            int c = 0;
            for (int j = 0; j < SE3_MAX_SAMPLE; j++)
                c += s[j] >= 0 ? 1 : 0;
            s_cnt = c;
        }
        pad_list(SE3_MAX_SAMPLE); // entries [count, 8) were not written: the sample list has at most 8
        if (tid < SE3_MAX_SAMPLE && s[tid] < 0) {
            g_w[tid] = 0.0f;
#pragma unroll
            for (int c = 0; c < 3; c++)
                g_f[c * capA + tid] = g_t[c * capA + tid] = 0.0f;
        }
    }
    __syncthreads();
    int cur = 0;
    if (warp != 1)
        group_tfc_fit<THREADS>(g_w, g_alpha, g_f, g_t, capA, s_cnt, sT[0], nullptr, tid); // getTransformFromMatches (:161-179)
    __syncthreads();

    DBG_STAMP(dbg, 17);
    bool aborted = false;
    for (int refinements = 1; refinements < 20; refinements++) {
        DBG_STAMP(dbg, 16 + refinements * 8);
        // computeInliersAndError (:181-214): score all matches under sT[cur]
        double Td[16];
#pragma unroll
        for (int i = 0; i < 16; i++)
            Td[i] = (double)sT[cur][i];
        double md_r[NPT];
        unsigned inbits = 0;
#pragma unroll
        for (int it = 0; it < NPT; it++) {
            const int word = it * NW + warp;
            md_r[it] = 0.0;
            if (word < words) {
                const int j = word * 32 + lane;
                bool in = false;
                if (j < M) {
                    const rgbdfe_point3f o = P1[j], t = P2[j];
                    if (!(o.z == 0.0f || t.x == 0.0f)) {
                        const double md = mahal2(o, t, Td, C, prm.rcx, prm.rcy);
                        in = !(md > th) && (md >= 0.0);
                        md_r[it] = md;
                    }
                }
                const unsigned bal = __ballot_sync(0xffffffffu, in);
                if (lane == 0) {
                    cur_m[word] = bal;
                    wbase[word] = __popc(bal);
                }
                inbits |= in ? (1u << it) : 0u;
            }
        }
        if (tid == THREADS - 1)
            s_abort = ld_volatile(&R.replay[pair].done);
        __syncthreads();
        DBG_STAMP(dbg, 16 + refinements * 8 + 1);
        const int abort_now = s_abort;
        if (warp == 0) { // exclusive prefix of the per-word counts
            int carry = 0;
            for (int b = 0; b < words; b += 32) {
                const int v = (b + lane < words) ? wbase[b + lane] : 0;
                int inc = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int n = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o)
                        inc += n;
                }
                if (b + lane < words)
                    wbase[b + lane] = carry + inc - v;
                carry += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) {
                s_cnt = carry;
                s_ctl = -1;
            }
        }
        __syncthreads();
        if (abort_now) { // the pair's loop already ended on an earlier iteration: this hypothesis is never looked at
            aborted = true;
            break;
        }
        const int cnt = s_cnt;
        // ordered inlier list: gather the inliers' data for the next fit and their d^2 for the error
#pragma unroll
        for (int it = 0; it < NPT; it++) {
            const int word = it * NW + warp;
            if (word < words && ((inbits >> it) & 1u)) {
                const int k = wbase[word] + __popc(cur_m[word] & ((1u << lane) - 1));
                g_d2[k] = md_r[it];
                gather(k, word * 32 + lane);
            }
        }
        pad_list(cnt);
        __syncthreads();
        DBG_STAMP(dbg, 16 + refinements * 8 + 2);
        if (dbg)
            dbg[16 + refinements * 8 + 7] = cnt;
        const bool more = refinements < 19 && cnt >= (int)prm.min_inlier_th;
        if (warp == 1) {
            if (lane == 0) {
                int stop = 0, accept = 0;
                if (cnt < (int)prm.min_inlier_th)
                    stop = 1;
                else {
                    double sum = 0.0; // FP64 sum in match order
                    int k = 0;
                    for (; k + 4 <= cnt; k += 4) {
                        const double a = g_d2[k], b = g_d2[k + 1], c = g_d2[k + 2], d = g_d2[k + 3];
                        sum += a;
                        sum += b;
                        sum += c;
                        sum += d;
                    }
                    for (; k < cnt; k++)
                        sum += g_d2[k];
                    double err = 1e9;
                    if (cnt >= 3)
                        err = sqrt(sum / (double)cnt);
                    if (err > (double)prm.max_mahal)
                        stop = 1;
                    else if (cnt >= s_refCount && err <= s_refErr) {
                        accept = 1;
                        if (cnt == s_refCount)
                            stop = 1;
                        s_refErr = err;
                        s_refCount = cnt;
                    } else
                        stop = 1;
                }
                *reinterpret_cast<volatile int*>(&s_ctl) = stop | (accept << 1);
            }
        } else if (more)
            group_tfc_fit<THREADS>(g_w, g_alpha, g_f, g_t, capA, cnt, sT[cur ^ 1], &s_ctl, tid, dbg ? dbg + 16 + refinements * 8 + 4 : nullptr); // next round's transform, speculatively
        __syncthreads();
        DBG_STAMP(dbg, 16 + refinements * 8 + 3);
        const int ctl = s_ctl;
        if (ctl & 2) { // vRefinedMatches = vInlierMatches; refinedTransformation = transformation
            unsigned* t = ref_m;
            ref_m = cur_m;
            cur_m = t;
            if (tid < 16)
                sRefT[tid] = sT[cur][tid];
        }
        if (ctl & 1)
            break;
        cur ^= 1;
    }
    if (aborted)
        return;
    __syncthreads();
    DBG_STAMP(dbg, 200);
    if (dbg)
        dbg[301] = clock64();
    HypOut* out = const_cast<HypOut*>(R.hyp_all) + (size_t)pair * MAX_ITERS + h;
    if (tid == 0) {
        out->count = s_refCount;
        out->err = s_refErr;
    }
    if (tid < 16)
        out->T[tid] = sRefT[tid];
    unsigned* om = const_cast<unsigned*>(R.mask_all) + ((size_t)pair * MAX_ITERS + h) * mw;
    for (int i = tid; i < words; i += THREADS)
        om[i] = ref_m[i];
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        *reinterpret_cast<volatile int*>(&R.ready_all[(size_t)pair * MAX_ITERS + h]) = 1;
        __threadfence();
    }
    if (warp == 0) {
        __syncwarp();
        warp_try_replay(R, pair, hyp_end, next_end, scratch, lane);
        DBG_STAMP(dbg, 201);
    }
}

// Grid: (CTAs per pair, pairs).  The first chunk of iterations gets one CTA per iteration; later chunks, which exist
// only for the pairs whose loop has not ended (mean executed iterations on tracking data: 1.1), get a few CTAs that
// walk the chunk's iterations in order and leave as soon as the pair is finished -- instead of one CTA per
// iteration launched just to find that out.
// MINB > 1 caps the registers per thread (65536 / (THREADS * MINB)): un-capped, a hypothesis CTA of any of these shapes takes the
// whole register file of its SM (512 x 128, 256 x 248), and a latency-bound CTA that mostly waits on dependent FP chains then keeps
// every throughput kernel off that SM.
template <int THREADS, int NPT, int MINB = 1>
__global__ void __launch_bounds__(THREADS, MINB) k_se3_hypotheses(ReplayCtx R, const float* __restrict__ w_all, int hyp_begin, int hyp_end, int next_end, int npairs)
{
    __shared__ int s_done, s_skip;
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    // The first chunk has one grid row per pair.  Later chunks exist only for the pairs whose loop has not ended (a few per
    // cent of a verification batch): a small grid walks the pairs, one uniform read each.
    for (int py = blockIdx.y; py < npairs; py += gridDim.y) {
        const int pair = R.slot0 + py;
        if (hyp_begin > 0) {
            if (threadIdx.x == 0) {
                const ReplayState* rp = R.replay + pair;
                s_skip = ld_volatile(&rp->done) == 2 ? 1 : 0; // finalised (also every pair that never ran)
            }
            __syncthreads();
            const int skip = s_skip;
            __syncthreads();
            if (skip)
                continue;
        }
        for (int h = hyp_begin + blockIdx.x; h < hyp_end || h == hyp_begin + (int)blockIdx.x; h += gridDim.x) {
            cta_se3_hypothesis<THREADS, NPT>(R, w_all, pair, h, hyp_begin, hyp_end, next_end, blockIdx.x == 0 && h == hyp_begin);
            const int hn = h + (int)gridDim.x;
            if (hn >= hyp_end)
                break;
            __syncthreads(); // the replay of warp 0 used the shared scratch
            if (threadIdx.x < 32) {
                // The next iteration of this CTA may not be drawn yet.  The replay draws more once everything drawn so far has
                // been consumed, but that needs every CTA of the pair to have run, and CTAs of a grid need not be co-resident
                // (small SM partitions, many contexts on one GPU): waiting for them could starve them of the slot this CTA
                // holds.  Samples depend only on the match count and the pair's stream, so this warp draws the chunk itself,
                // under the pair lock, whenever nobody has yet.
                const ReplayState* rp = R.replay + pair;
                int d;
                while (!(d = ld_volatile(&rp->done) | (ld_volatile(&rp->run) ^ 1)) && next_end < 0 && ld_volatile(&rp->sampled_k) <= hn)
                    warp_draw_ahead(R, pair, hn, -next_end, reinterpret_cast<unsigned*>(smem_dyn), threadIdx.x);
                if (threadIdx.x == 0)
                    s_done = d;
            }
            __syncthreads();
            if (s_done)
                break;
        }
        __syncthreads(); // the next pair reuses the shared state
    }
}

template <int THREADS, int NPT, int MINB = 1>
static int launch_hyp(rgbdfe_ctx* c, const ReplayCtx& R, int npairs, size_t smem, int b, int e, int next_end, int max_ctas)
{
    CUDA_TRY(c, (ensure_dyn_smem(c, k_se3_hypotheses<THREADS, NPT, MINB>, smem)));
    int nblk = e - b > 0 ? e - b : 1; // at least one CTA per pair: it drives the replay when there is nothing to evaluate
    if (nblk > max_ctas)
        nblk = max_ctas;
    const int rows = b == 0 ? npairs : (npairs < 296 ? npairs : 296); // later chunks: at most 8 x 296 CTAs walk the pairs (an all-finished batch costs a few idle waves, a hard one keeps every SM busy)
    k_se3_hypotheses<THREADS, NPT, MINB><<<dim3(nblk, rows), THREADS, smem, c->stream>>>(R, c->d_w, b, e, next_end, npairs);
    c->launches++;
    return RGBDFE_OK;
}

int launch_ransac_se3(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, const Se3Params& prm, int update_f2, int per_pair_rng,
    unsigned seed, int first_pair_index, int min_matches, rgbdfe_se3_result* d_results, const FusedFilter* ff)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    if (prm.sample_size > SE3_MAX_SAMPLE || prm.sample_size < 1) {
        c->set_error("sample_size must be in [1, 8]");
        return RGBDFE_E_ARG;
    }
    const int cap = c->kp_cap;
    if (cap > 4096) {
        c->set_error("RANSAC: more than 4096 matches per pair are not supported");
        return RGBDFE_E_CAPACITY;
    }
    const int iters = prm.iterations < MAX_ITERS ? (prm.iterations > 0 ? prm.iterations : 0) : MAX_ITERS;
    // chunk boundaries: a single pair fills the GPU with hypotheses; many pairs start small
    const bool few = npairs < 16;
    const int b1 = few ? c->ransac_first_chunk : 8, b2 = few ? (iters > b1 ? iters : b1 + 1) : 64; // few pairs: one follow-up launch takes the rest
    const size_t prep_smem = (size_t)cap * (sizeof(SKey) + 3 * sizeof(int)) + 2 * SORT_QCAP * sizeof(int3);
    const int capA = ((cap + 3) & ~3) + 8;
    size_t hyp_smem = (size_t)capA * 40 + (size_t)c->mask_words * 12;
    const size_t min_smem = ((size_t)RNG_SCRATCH_WORDS + c->mask_words) * 4;
    if (hyp_smem < min_smem)
        hyp_smem = min_smem;
    CUDA_TRY(c, ensure_dyn_smem(c, k_se3_prepare, prep_smem));
    {
        ProfScope prof(c, KID_SE3_PREPARE);
        FusedFilter f = { 0, 0.0f, 0, 0, 0 };
        if (ff)
            f = *ff;
        k_se3_prepare<<<npairs, PREP_THREADS, prep_smem, c->stream>>>(d_pairs, slot0, f, c->d_partial, c->d_matches, c->d_sorted, c->d_P1, c->d_P2, c->d_w, cap, c->d_nmatch,
            prm, update_f2, per_pair_rng, seed, first_pair_index, min_matches, c->d_rng, c->d_pair_rng, c->d_latch, c->d_samples, c->d_rng_used,
            c->d_replay, c->d_hyp_ready, c->d_pair_lock, b1, c->dbg_on ? c->d_dbg : nullptr);
        c->launches++;
    }
    ReplayCtx R;
    R.pairs = d_pairs;
    R.sorted_all = c->d_sorted;
    R.P1_all = c->d_P1;
    R.P2_all = c->d_P2;
    R.match_cap = cap;
    R.prm = prm;
    R.update_f2 = update_f2;
    R.per_pair_rng = per_pair_rng;
    R.ctx_rng = c->d_rng;
    R.pair_rng = c->d_pair_rng;
    R.latch = c->d_latch;
    R.samples_all = c->d_samples;
    R.rng_used_all = c->d_rng_used;
    R.hyp_all = c->d_hyp;
    R.mask_all = c->d_hyp_mask;
    R.mask_words = c->mask_words;
    R.replay = c->d_replay;
    R.ready_all = c->d_hyp_ready;
    R.lock_all = c->d_pair_lock;
    R.results = d_results;
    R.slot0 = slot0;
    R.dbg = c->dbg_on ? c->d_dbg : nullptr;
    R.inl_pos_all = c->d_inl_pos;
    // Few pairs (tracking chains): ONE launch of b1 CTAs per pair.  CTA k evaluates iterations k, k + b1, ... until the
    // pair's loop has ended; iterations beyond the first b1 are drawn b1 at a time by the replay when needed.
    // Many pairs (verification): chunks of 8, 56 and the rest, the later ones on 8 walking CTAs per pair.
    const int bounds[4] = { 0, few ? iters : b1, few ? iters : b2, iters };
    for (int s = 0; s < 3; s++) {
        const int b = bounds[s] < iters ? bounds[s] : iters;
        const int e = (s == 2 || bounds[s + 1] > iters) ? iters : bounds[s + 1];
        const int next_end = few ? -b1 : (s == 0 ? (b2 < iters ? b2 : iters) : iters);
        const int max_ctas = few ? b1 : (s == 0 ? b1 : 8);
        if (s > 0 && e <= b)
            break;
        ProfScope prof(c, KID_SE3_HYP);
        int rc;
        // Tracking chains keep 512-thread CTAs without a register cap: capped 128- / 256-thread CTAs (128 registers,
        // ~180 B of spills) measured 74-77 k frames/s against 78-79 k on the headline workload.
        if (few && c->ransac_small_ctas && cap <= 2304) {
            if (cap <= 1152)
                rc = launch_hyp<128, 9, 4>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
            else
                rc = launch_hyp<256, 9, 2>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
        } else if (few || cap > 2304) {
            if (cap <= 1536)
                rc = launch_hyp<512, 3>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
            else if (cap <= 2560)
                rc = launch_hyp<512, 5>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
            else
                rc = launch_hyp<512, 8>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
        } else if (c->ransac_verify_cap == 0 || (c->ransac_verify_cap < 0 && npairs < -c->ransac_verify_cap)) { // (A/B knob)
            rc = cap <= 1152 ? launch_hyp<128, 9>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas) : launch_hyp<256, 9>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
        } else if (cap <= 1152) // verification batches: thousands of CTAs, occupancy matters more than the spills
            rc = launch_hyp<128, 9, 4>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
        else
            rc = launch_hyp<256, 9, 2>(c, R, npairs, hyp_smem, b, e, next_end, max_ctas);
        if (rc != RGBDFE_OK)
            return rc;
        if (e >= iters)
            break;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
