// K9-K11 — RansacSE3::compute (Solver/SolverSE3.cpp:23-133) for a batch of pairs.
//
// The reference loop is sequential only in (a) the rand() stream and (b) the accept /
// "n += 10" / break control; each iteration's refinement result depends on nothing but
// its sample.  So:
//   k_se3_prepare    1 CTA / pair: warp 0 sorts the matches exactly as libstdc++ std::sort
//                    does (introsort.cuh) and latches the depth covariance; warp 1 draws
//                    the biased 4-samples of ALL iterations from the glibc-compatible stream
//                    (samples depend only on M, never on the data), recording the stream
//                    position after every iteration.
//   k_se3_hypotheses 1 warp / (pair, iteration): up to 19 refinement rounds of
//                    {TFC fit (FP32) -> Mahalanobis scoring of all matches (FP64, lanes in
//                    parallel, ballot = inlier set, ordered FP64 error sum)}.
//   k_se3_replay     1 warp / pair: replays SolverSE3.cpp:86-101 over the hypothesis results
//                    to find which iterations the reference would have executed, the best
//                    model, the identity fallback (:105-117), the Frame flag updates
//                    (:38-42,119-123) and rewinds the RNG stream to the right position.
// Hypotheses are evaluated in growing chunks (8, 56, rest); pairs whose loop already
// terminated skip later chunks, so the common early exit (> 80 % inliers) costs one chunk.
#include "introsort.cuh"
#include "se3_math.cuh"

namespace rgbdfe {

#define SE3_MAX_SAMPLE 8
#define HYP_WARPS 8

__device__ __forceinline__ int rng_next(RngState& s)
{
    unsigned v = (unsigned)s.r[(s.idx - 31) % 34] + (unsigned)s.r[(s.idx - 3) % 34];
    s.r[s.idx % 34] = (int)v;
    s.idx++;
    if (s.idx >= 34 * 1000000)
        s.idx -= 34 * 999990;
    return (int)(v >> 1);
}

__device__ void rng_seed(RngState& s, unsigned seed)
{
    // glibc srandom_r, TYPE_3: r[i] = 16807 * r[i-1] % 2147483647, then 310 discarded outputs
    unsigned lin[34];
    if (seed == 0)
        seed = 1;
    lin[0] = seed;
    for (int i = 1; i < 31; i++) {
        long long w = (16807LL * (int)lin[i - 1]) % 2147483647;
        if (w < 0)
            w += 2147483647;
        lin[i] = (unsigned)w;
    }
    for (int i = 31; i < 34; i++)
        lin[i] = lin[i - 31];
    for (int i = 0; i < 34; i++)
        s.r[i] = (int)lin[i];
    s.idx = 34;
    for (int i = 34; i < 344; i++)
        rng_next(s);
}

// Random::randomInt (System/Random.cpp:15-20)
__device__ __forceinline__ int random_int(RngState& s, int mn, int mx)
{
    int d = mx - mn + 1;
    return int(((double)rng_next(s) / ((double)2147483647 + 1.0)) * d) + mn;
}

__global__ void __launch_bounds__(64) k_se3_prepare(const MatchPair* __restrict__ pairs, rgbdfe_dmatch* __restrict__ matches_all, int match_cap,
    const int* __restrict__ nmatch, Se3Params prm, int update_f2, int per_pair_rng, unsigned seed, int first_pair_index, int min_matches,
    RngState* __restrict__ ctx_rng, RngState* __restrict__ pair_rng, DepthLatch* __restrict__ latch, int* __restrict__ samples_all,
    int* __restrict__ rng_used_all, ReplayState* __restrict__ replay, int n_hyp_max)
{
    const int pair = blockIdx.x;
    const MatchPair P = pairs[pair];
    rgbdfe_dmatch* m = matches_all + (size_t)pair * match_cap;
    const int M = nmatch[pair];
    const int tid = threadIdx.x;
    const bool run = M >= (int)prm.min_inlier_th && M >= min_matches; // SolverSE3.cpp:29 / PoseGraph.cpp:142-144
    if (tid == 0) {
        ReplayState st;
        st.n = 0;
        st.k = 0;
        st.done = run ? 0 : 1;
        st.validIters = 0;
        st.bestHyp = -1;
        st.bestCount = 0;
        st.rmse = 1e6f;
        st.run = run ? 1 : 0;
        st.nm = M;
        st.rng_calls = 0;
        replay[pair] = st;
    }
    if (!run)
        return;
    if (tid == 0) {
        // SolverSE3.cpp:38-42: every matched train keypoint of F2 becomes an outlier
        if (update_f2 && P.t_outlier)
            for (int i = 0; i < M; i++)
                P.t_outlier[m[i].trainIdx] = 1;
        std_sort(m, m + M, DmLess()); // :52
        // depthCovariance latch (:282-287): the first source depth that reaches errorFunction2
        if (!latch->latched) {
            for (int i = 0; i < M; i++) {
                const rgbdfe_point3f o = P.q_xyz[m[i].queryIdx];
                const rgbdfe_point3f t = P.t_xyz[m[i].trainIdx];
                if (o.z == 0.0f || t.x == 0.0f)
                    continue;
                if (isnan(o.z) || isnan(t.z))
                    continue;
                double z = (double)o.z;
                double sd = 0.01 * z * z;
                latch->cov = sd * sd;
                __threadfence();
                latch->latched = 1;
                break;
            }
        }
    }
    if (tid == 32) {
        // sampleMatches (:135-159) for every iteration the loop could execute
        RngState s;
        if (per_pair_rng) {
            rng_seed(s, seed + (unsigned)(first_pair_index + pair));
            pair_rng[pair] = s;
        } else
            s = *ctx_rng;
        int* samples = samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE;
        int* used = rng_used_all + (size_t)pair * MAX_ITERS;
        const int ss = (int)prm.sample_size;
        int calls = 0;
        const int iters = (M >= ss) ? min(prm.iterations, n_hyp_max) : 0;
        for (int k = 0; k < iters; k++) {
            int ids[SE3_MAX_SAMPLE];
            int cnt = 0, safetyNet = 0;
            while (cnt < ss) {
                int id1 = random_int(s, 0, M - 1);
                int id2 = random_int(s, 0, M - 1);
                calls += 2;
                if (id1 > id2)
                    id1 = id2;
                // std::set insert: keep ascending, ignore duplicates
                int pos = 0;
                while (pos < cnt && ids[pos] < id1)
                    pos++;
                if (!(pos < cnt && ids[pos] == id1)) {
                    for (int j = cnt; j > pos; j--)
                        ids[j] = ids[j - 1];
                    ids[pos] = id1;
                    cnt++;
                }
                if (++safetyNet > 10000)
                    break;
            }
            for (int j = 0; j < SE3_MAX_SAMPLE; j++)
                samples[k * SE3_MAX_SAMPLE + j] = j < cnt ? ids[j] : -1;
            used[k] = calls;
        }
    }
}

// Scores every match against T (computeInliersAndError, SolverSE3.cpp:181-214), one warp.
// Writes the inlier set as a bitmask over sorted-match positions, returns count and the
// mean error.  The FP64 error sum runs in match order, as the reference's loop does.
__device__ __forceinline__ double warp_score(const rgbdfe_point3f* __restrict__ P1, const rgbdfe_point3f* __restrict__ P2, int M,
    const float* T, double th, double C, double rcx, double rcy, unsigned* mask, int lane, int* out_count)
{
    double Td[16];
#pragma unroll
    for (int i = 0; i < 16; i++)
        Td[i] = (double)T[i];
    double sum = 0.0;
    int cnt = 0;
    const int words = (M + 31) >> 5;
    for (int w = 0; w < words; w++) {
        const int j = w * 32 + lane;
        double d2 = 0.0;
        bool in = false;
        if (j < M) {
            const rgbdfe_point3f o = P1[j], t = P2[j];
            if (!(o.z == 0.0f || t.x == 0.0f)) {
                const double md = mahal2(o, t, Td, C, rcx, rcy);
                if (!(md > th) && (md >= 0.0)) {
                    in = true;
                    d2 = md;
                }
            }
        }
        unsigned bal = __ballot_sync(0xffffffffu, in);
        if (lane == 0)
            mask[w] = bal;
        cnt += __popc(bal);
        while (bal) {
            const int l = __ffs(bal) - 1;
            bal &= bal - 1;
            sum += __shfl_sync(0xffffffffu, d2, l);
        }
    }
    __syncwarp();
    *out_count = cnt;
    if (cnt < 3)
        return 1e9;
    double mean = sum / (double)cnt;
    return sqrt(mean);
}

// dynamic smem: rgbdfe_point3f P1[match_cap], P2[match_cap]; unsigned masks[HYP_WARPS][2][mask_words]
__global__ void __launch_bounds__(HYP_WARPS * 32) k_se3_hypotheses(const MatchPair* __restrict__ pairs, const rgbdfe_dmatch* __restrict__ matches_all,
    int match_cap, const int* __restrict__ nmatch, Se3Params prm, const DepthLatch* __restrict__ latch, const int* __restrict__ samples_all,
    HypOut* __restrict__ hyp_all, unsigned* __restrict__ mask_all, int mask_words, const ReplayState* __restrict__ replay, int hyp_begin,
    int hyp_end)
{
    const int pair = blockIdx.y;
    const ReplayState st = replay[pair];
    if (st.done || !st.run)
        return;
    const int M = nmatch[pair];
    if (M < (int)prm.sample_size)
        return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    rgbdfe_point3f* P1 = reinterpret_cast<rgbdfe_point3f*>(smem_raw);
    rgbdfe_point3f* P2 = P1 + match_cap;
    unsigned* masks = reinterpret_cast<unsigned*>(P2 + match_cap);
    const MatchPair P = pairs[pair];
    const rgbdfe_dmatch* m = matches_all + (size_t)pair * match_cap;
    for (int j = threadIdx.x; j < M; j += blockDim.x) {
        const rgbdfe_dmatch mm = m[j];
        P1[j] = P.q_xyz[mm.queryIdx];
        P2[j] = P.t_xyz[mm.trainIdx];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = hyp_begin + blockIdx.x * HYP_WARPS + warp;
    if (h >= hyp_end || h >= prm.iterations || h >= MAX_ITERS)
        return;
    unsigned* cur = masks + (size_t)warp * 2 * mask_words;
    unsigned* refined = cur + mask_words;
    const int words = (M + 31) >> 5;
    for (int w = lane; w < words; w += 32)
        cur[w] = 0;
    __syncwarp();
    if (lane == 0) {
        const int* s = samples_all + ((size_t)pair * MAX_ITERS + h) * SE3_MAX_SAMPLE;
        for (int j = 0; j < SE3_MAX_SAMPLE; j++)
            if (s[j] >= 0)
                cur[s[j] >> 5] |= 1u << (s[j] & 31);
    }
    __syncwarp();
    const double C = latch->cov;
    const double th = (double)(prm.max_mahal * prm.max_mahal);
    int refinedCount = 0;
    double refinedErr = 1e6;
    float refinedT[16];
#pragma unroll
    for (int i = 0; i < 16; i++)
        refinedT[i] = (i % 5 == 0) ? 1.0f : 0.0f;

    for (int refinements = 1; refinements < 20; refinements++) {
        // getTransformFromMatches over the current inliers, in match order (all lanes redundantly)
        Tfc tfc;
        tfc.reset();
        for (int w = 0; w < words; w++) {
            unsigned bits = cur[w];
            while (bits) {
                const int j = w * 32 + __ffs(bits) - 1;
                bits &= bits - 1;
                tfc.add(P1[j], P2[j]);
            }
        }
        float T[16];
        tfc.transform(T);
        __syncwarp();
        int cnt;
        const double err = warp_score(P1, P2, M, T, th, C, prm.rcx, prm.rcy, cur, lane, &cnt);
        if (cnt < (int)prm.min_inlier_th || err > (double)prm.max_mahal)
            break;
        if (cnt >= refinedCount && err <= refinedErr) {
            const int prev = refinedCount;
#pragma unroll
            for (int i = 0; i < 16; i++)
                refinedT[i] = T[i];
            for (int w = lane; w < words; w += 32)
                refined[w] = cur[w];
            __syncwarp();
            refinedErr = err;
            refinedCount = cnt;
            if (cnt == prev)
                break;
        } else
            break;
    }
    HypOut* out = hyp_all + (size_t)pair * MAX_ITERS + h;
    if (lane == 0) {
        out->count = refinedCount;
        out->err = refinedErr;
#pragma unroll
        for (int i = 0; i < 16; i++)
            out->T[i] = refinedT[i];
    }
    unsigned* om = mask_all + ((size_t)pair * MAX_ITERS + h) * mask_words;
    for (int w = lane; w < words; w += 32)
        om[w] = refinedCount > 0 ? refined[w] : 0u;
}

// One warp per pair.  dynamic smem: unsigned mask[mask_words] (identity fallback)
__global__ void __launch_bounds__(32) k_se3_replay(const MatchPair* __restrict__ pairs, const rgbdfe_dmatch* __restrict__ matches_all, int match_cap,
    const int* __restrict__ nmatch, Se3Params prm, int update_f2, int per_pair_rng, RngState* __restrict__ ctx_rng,
    const DepthLatch* __restrict__ latch, const int* __restrict__ rng_used_all, const HypOut* __restrict__ hyp_all,
    const unsigned* __restrict__ mask_all, int mask_words, ReplayState* __restrict__ replay, int hyp_end, int final_stage,
    rgbdfe_se3_result* __restrict__ results, int* __restrict__ inl_pos_all)
{
    extern __shared__ unsigned s_mask[];
    const int pair = blockIdx.x, lane = threadIdx.x;
    ReplayState st = replay[pair];
    if (st.done == 2)
        return; // already finalised in an earlier stage
    const int M = st.nm;
    const MatchPair P = pairs[pair];
    const rgbdfe_dmatch* m = matches_all + (size_t)pair * match_cap;
    rgbdfe_se3_result* res = results + pair;
    int* inl_pos = inl_pos_all + (size_t)pair * match_cap;

    if (!st.run) {
        if (lane == 0) {
            res->ok = 0;
            res->rmse = 1e6f;
            for (int i = 0; i < 16; i++)
                res->T21[i] = (i % 5 == 0) ? 1.0f : 0.0f;
            res->n_inliers = 0;
            res->iters_executed = 0;
            res->n_matches = M;
            st.done = 2;
            replay[pair] = st;
        }
        return;
    }
    const bool Mok = M >= (int)prm.sample_size;
    if (!st.done) {
        // all lanes replay redundantly (uniform control flow)
        bool broke = false;
        const HypOut* hyp = hyp_all + (size_t)pair * MAX_ITERS;
        while (st.n < prm.iterations && Mok && st.k < hyp_end && st.k < MAX_ITERS) {
            const int cnt = hyp[st.k].count;
            const double err = hyp[st.k].err;
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
        if (broke || st.n >= prm.iterations || !Mok || st.k >= MAX_ITERS)
            st.done = 1;
    }
    if (!st.done) {
        if (lane == 0)
            replay[pair] = st;
        return; // needs the next chunk of hypotheses
    }
    (void)final_stage;
    // ---- finalise ----
    st.rng_calls = st.k > 0 ? rng_used_all[(size_t)pair * MAX_ITERS + st.k - 1] : 0;
    if (!per_pair_rng && lane == 0) {
        RngState s = *ctx_rng;
        for (int i = 0; i < st.rng_calls; i++)
            rng_next(s);
        *ctx_rng = s;
    }
    float T[16];
#pragma unroll
    for (int i = 0; i < 16; i++)
        T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    int n_inl = 0;
    float rmse = st.rmse;
    const unsigned* mask = nullptr;
    if (st.validIters == 0) {
        // identity fallback (SolverSE3.cpp:105-117)
        int cnt;
        const double th = (double)(prm.max_mahal * prm.max_mahal);
        // gather points on the fly: P1/P2 views are indirect here, so score through a small adaptor
        double sum = 0.0;
        cnt = 0;
        double Td[16];
#pragma unroll
        for (int i = 0; i < 16; i++)
            Td[i] = (double)T[i];
        const int words = (M + 31) >> 5;
        const double C = latch->cov;
        for (int w = 0; w < words; w++) {
            const int j = w * 32 + lane;
            double d2 = 0.0;
            bool in = false;
            if (j < M) {
                const rgbdfe_point3f o = P.q_xyz[m[j].queryIdx], t = P.t_xyz[m[j].trainIdx];
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
        const HypOut* h = hyp_all + (size_t)pair * MAX_ITERS + st.bestHyp;
#pragma unroll
        for (int i = 0; i < 16; i++)
            T[i] = h->T[i];
        n_inl = st.bestCount;
        mask = mask_all + ((size_t)pair * MAX_ITERS + st.bestHyp) * mask_words;
    }
    const int ok = n_inl >= (int)prm.min_inlier_th ? 1 : 0;
    // inlier positions (sorted-match order) + Frame::setInlier (:119-123)
    if (mask && n_inl > 0) {
        int base = 0;
        const int words = (M + 31) >> 5;
        for (int w = 0; w < words; w++) {
            const unsigned bits = mask[w];
            if (bits & (1u << lane)) {
                const int j = w * 32 + lane;
                inl_pos[base + __popc(bits & ((1u << lane) - 1))] = j;
                if (ok && update_f2 && P.t_outlier)
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
        st.done = 2;
        replay[pair] = st;
    }
}

int launch_ransac_se3(rgbdfe_ctx* c, const MatchPair* d_pairs, int npairs, const Se3Params& prm, int update_f2, int per_pair_rng, unsigned seed,
    int first_pair_index, int min_matches, rgbdfe_se3_result* d_results)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    if (prm.sample_size > SE3_MAX_SAMPLE || prm.sample_size < 1) {
        c->set_error("sample_size must be in [1, 8]");
        return RGBDFE_E_ARG;
    }
    const int iters = prm.iterations < MAX_ITERS ? prm.iterations : MAX_ITERS;
    {
    ProfScope prof(c, KID_SE3_PREPARE);
    k_se3_prepare<<<npairs, 64, 0, c->stream>>>(d_pairs, c->d_matches, c->kp_cap, c->d_nmatch, prm, update_f2, per_pair_rng, seed, first_pair_index,
        min_matches, c->d_rng, c->d_pair_rng, c->d_latch, c->d_samples, c->d_rng_used, c->d_replay, MAX_ITERS);
    c->launches++;
    }
    const size_t hyp_smem = (size_t)c->kp_cap * 24 + (size_t)HYP_WARPS * 2 * c->mask_words * 4;
    static size_t configured = 0;
    if (hyp_smem > configured) {
        CUDA_TRY(c, cudaFuncSetAttribute(k_se3_hypotheses, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hyp_smem));
        configured = hyp_smem;
    }
    const int bounds[4] = { 0, 8, 64, iters };
    bool any = false;
    for (int s = 0; s < 3; s++) {
        int b = bounds[s], e = bounds[s + 1] < iters ? bounds[s + 1] : iters;
        if (s == 2)
            e = iters;
        if (e <= b && any)
            continue;
        const bool last = e >= iters;
        if (e > b) {
            ProfScope prof(c, KID_SE3_HYP);
            k_se3_hypotheses<<<dim3((e - b + HYP_WARPS - 1) / HYP_WARPS, npairs), HYP_WARPS * 32, hyp_smem, c->stream>>>(d_pairs, c->d_matches,
                c->kp_cap, c->d_nmatch, prm, c->d_latch, c->d_samples, c->d_hyp, c->d_hyp_mask, c->mask_words, c->d_replay, b, e);
            c->launches++;
        }
        ProfScope prof2(c, KID_SE3_REPLAY);
        k_se3_replay<<<npairs, 32, (size_t)c->mask_words * 4, c->stream>>>(d_pairs, c->d_matches, c->kp_cap, c->d_nmatch, prm, update_f2, per_pair_rng,
            c->d_rng, c->d_latch, c->d_rng_used, c->d_hyp, c->d_hyp_mask, c->mask_words, c->d_replay, e, last ? 1 : 0, d_results, c->d_inl_pos);
        c->launches++;
        any = true;
        if (last)
            break;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
