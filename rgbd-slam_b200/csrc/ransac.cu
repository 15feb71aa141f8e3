// K9-K11 — RansacSE3::compute (Solver/SolverSE3.cpp:23-133) for a batch of pairs.
//
// The reference loop is sequential only in (a) the rand() stream and (b) the accept /
// "n += 10" / break control; each iteration's refinement result depends on nothing but
// its sample.  So:
//   k_se3_prepare    1 CTA / pair: sorts the matches into exactly the permutation libstdc++'s
//                    std::sort produces (parallel introsort, see cta_std_sort), gathers the
//                    matched 3D points and TFC weights in sorted order, latches the depth
//                    covariance, marks F2 outliers, and draws the biased 4-samples of the first
//                    chunk of iterations from the glibc-compatible stream (samples depend only
//                    on M, never on the data).
//   k_se3_hypotheses 1 CTA / (pair, iteration): up to 19 refinement rounds of
//                    {TFC fit (FP32, 9 covariance lanes) -> Mahalanobis scoring of all matches
//                    (FP64, all threads) -> ordered inlier list -> ordered FP64 error sum}.
//   k_se3_replay     1 warp / pair: replays SolverSE3.cpp:86-101 over the hypothesis results
//                    to find which iterations the reference would have executed, the best
//                    model, the identity fallback (:105-117), the Frame flag updates
//                    (:119-123), rewinds the RNG stream to the right position, and samples
//                    the next chunk when the loop has not ended yet.
// Hypotheses are evaluated in growing chunks; pairs whose loop already terminated skip later
// chunks, so the common early exit costs one chunk.
#include "introsort.cuh"
#include "se3_math.cuh"

namespace rgbdfe {

#define SE3_MAX_SAMPLE 8
#define PREP_THREADS 256
#define HYP_THREADS 128

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

// sampleMatches (SolverSE3.cpp:135-159) for iterations [k0, k1); one thread.
__device__ void sample_chunk(RngState& s, int M, int ss, int k0, int k1, int* __restrict__ samples, int* __restrict__ used, int& calls)
{
    for (int k = k0; k < k1; k++) {
        int ids[SE3_MAX_SAMPLE];
        int cnt = 0, safetyNet = 0;
        while (cnt < ss) {
            int id1 = random_int(s, 0, M - 1);
            int id2 = random_int(s, 0, M - 1);
            calls += 2;
            if (id1 > id2)
                id1 = id2;
            int pos = 0; // std::set insert: ascending, duplicates ignored
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

// ---------------------------------------------------------------------------------------
// Parallel std::sort.  libstdc++'s introsort is reproduced range by range: a warp partitions
// one range with the parallel form of __unguarded_partition — L-stoppers (a >= pivot) in
// ascending position and R-stoppers (a <= pivot) in descending position are compacted with
// ballots; the k-th swap of the sequential loop exchanges the k-th of each list while
// L[k] < R[k]; the returned cut is min(L[K], R[K-1]) — which yields the same array as the
// sequential scan (checked on the CPU against std::sort, tests/cpp/test_introsort.cpp).
// Ranges of one recursion level are independent, so levels are processed breadth-first by
// all warps.  Ranges <= 16 are finished by the final insertion sort, which never moves an
// element across a range boundary, so each leaf is insertion-sorted by one thread.
// ---------------------------------------------------------------------------------------
struct SKey {
    float d;
    int i;
};
struct KeyLess {
    __device__ bool operator()(const SKey& a, const SKey& b) const { return a.d < b.d; }
};

__device__ void warp_partition_range(SKey* a, int f, int l, int* Lpos, int* Rpos, int lane, int* cut_out)
{
    if (lane == 0) {
        int mid = f + (l - f) / 2;
        move_median_to_first_(a + f, a + f + 1, a + mid, a + l - 1, KeyLess());
    }
    __syncwarp();
    const float p = a[f].d;
    int nL = 0, nR = 0;
    const int first = f + 1;
    const unsigned lt = (1u << lane) - 1;
    for (int x0 = first; x0 < l; x0 += 32) {
        const int x = x0 + lane;
        const bool isL = x < l && !(a[x].d < p);
        const unsigned bal = __ballot_sync(0xffffffffu, isL);
        if (isL)
            Lpos[f + nL + __popc(bal & lt)] = x;
        nL += __popc(bal);
    }
    for (int x0 = l - 1; x0 >= first; x0 -= 32) {
        const int x = x0 - lane;
        const bool isR = x >= first && !(p < a[x].d);
        const unsigned bal = __ballot_sync(0xffffffffu, isR);
        if (isR)
            Rpos[f + nR + __popc(bal & lt)] = x;
        nR += __popc(bal);
    }
    __syncwarp();
    const int nmin = min(nL, nR);
    int K = nmin;
    for (int k0 = 0; k0 < nmin; k0 += 32) {
        const int k = k0 + lane;
        const bool stop = k < nmin && !(Lpos[f + k] < Rpos[f + k]);
        const unsigned bal = __ballot_sync(0xffffffffu, stop);
        if (bal) {
            K = k0 + __ffs(bal) - 1;
            break;
        }
    }
    const int Lk = K < nL ? Lpos[f + K] : l;
    const int Rprev = K > 0 ? Rpos[f + K - 1] : l;
    const int cut = min(Lk, Rprev);
    __syncwarp();
    for (int k = lane; k < K; k += 32) {
        const int x = Lpos[f + k], y = Rpos[f + k];
        const SKey t = a[x];
        a[x] = a[y];
        a[y] = t;
    }
    __syncwarp();
    *cut_out = cut;
}

// All PREP_THREADS threads call this.  a: smem keys; scratch: Lpos[n], Rpos[n], leaves[n],
// lists: 2 x QCAP ranges (f, l, depth packed in int3).
#define SORT_QCAP 512
// Called by the first `nthr` threads of the CTA (a multiple of 32); synchronises on named barrier 1.
#define SORT_BAR() asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory")
__device__ void cta_std_sort(SKey* a, int n, int* Lpos, int* Rpos, int* leaves, int3* qa, int3* qb, int* counters /* [3] */, int nthr)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
    if (n <= 1)
        return;
    if (tid == 0) {
        int lg = 0;
        for (int t = n; t > 1; t >>= 1)
            lg++;
        counters[0] = 0; // next-list size
        counters[1] = 0; // leaf count
        if (n > 16) {
            qa[0] = make_int3(0, n, 2 * lg);
            counters[2] = 1; // current-list size
        } else {
            leaves[0] = 0 | (n << 16);
            counters[1] = 1;
            counters[2] = 0;
        }
    }
    SORT_BAR();
    int3* cur = qa;
    int3* nxt = qb;
    while (true) {
        const int ncur = counters[2];
        if (ncur == 0)
            break;
        for (int r = warp; r < ncur; r += nwarps) {
            const int3 rg = cur[r];
            const int f = rg.x, l = rg.y, depth = rg.z;
            if (depth == 0) { // heapsort fallback of __introsort_loop; the range is then fully sorted
                if (lane == 0)
                    heap_sort_(a + f, a + l, KeyLess());
                __syncwarp();
                continue;
            }
            int cut;
            warp_partition_range(a, f, l, Lpos, Rpos, lane, &cut);
            if (lane == 0) {
                // right part [cut, l) and left part [f, cut), both with depth - 1
                const int parts[2][2] = { { cut, l }, { f, cut } };
                for (int s = 0; s < 2; s++) {
                    const int pf = parts[s][0], pl = parts[s][1];
                    if (pl - pf > 16) {
                        const int slot = atomicAdd(&counters[0], 1);
                        if (slot < SORT_QCAP)
                            nxt[slot] = make_int3(pf, pl, depth - 1);
                    } else if (pl - pf > 1) {
                        const int slot = atomicAdd(&counters[1], 1);
                        leaves[slot] = pf | (pl << 16);
                    }
                }
            }
        }
        SORT_BAR();
        if (tid == 0) {
            counters[2] = min(counters[0], SORT_QCAP);
            counters[0] = 0;
        }
        int3* t = cur;
        cur = nxt;
        nxt = t;
        SORT_BAR();
    }
    // final insertion sort, leaf by leaf (stable insertion == __final_insertion_sort restricted to the leaf)
    const int nleaves = counters[1];
    for (int r = tid; r < nleaves; r += nthr) {
        const int f = leaves[r] & 0xffff, l = leaves[r] >> 16;
        for (int i = f + 1; i < l; i++) {
            const SKey v = a[i];
            int j = i - 1;
            while (j >= f && v.d < a[j].d) {
                a[j + 1] = a[j];
                j--;
            }
            a[j + 1] = v;
        }
    }
    SORT_BAR();
}

// dynamic smem: SKey keys[cap]; int Lpos[cap], Rpos[cap], leaves[cap]; int3 qa[QCAP], qb[QCAP]
__global__ void __launch_bounds__(PREP_THREADS) k_se3_prepare(const MatchPair* __restrict__ pairs, const rgbdfe_dmatch* __restrict__ matches_all,
    rgbdfe_dmatch* __restrict__ sorted_all, rgbdfe_point3f* __restrict__ P1_all, rgbdfe_point3f* __restrict__ P2_all, float* __restrict__ w_all,
    int match_cap, const int* __restrict__ nmatch, Se3Params prm, int update_f2, int per_pair_rng, unsigned seed, int first_pair_index,
    int min_matches, const RngState* __restrict__ ctx_rng, RngState* __restrict__ pair_rng, DepthLatch* __restrict__ latch,
    int* __restrict__ samples_all, int* __restrict__ rng_used_all, ReplayState* __restrict__ replay, int first_chunk)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SKey* keys = reinterpret_cast<SKey*>(smem_raw);
    int* Lpos = reinterpret_cast<int*>(keys + match_cap);
    int* Rpos = Lpos + match_cap;
    int* leaves = Rpos + match_cap;
    int3* qa = reinterpret_cast<int3*>(leaves + match_cap);
    int3* qb = qa + SORT_QCAP;
    __shared__ int s_counters[4];

    const int pair = blockIdx.x;
    const MatchPair P = pairs[pair];
    const rgbdfe_dmatch* m = matches_all + (size_t)pair * match_cap;
    rgbdfe_dmatch* sorted = sorted_all + (size_t)pair * match_cap;
    const int M = nmatch[pair];
    const int tid = threadIdx.x;
    const bool run = M >= (int)prm.min_inlier_th && M >= min_matches; // SolverSE3.cpp:29 / PoseGraph.cpp:142-144
    const int ss = (int)prm.sample_size;
    const int n_first = (M >= ss) ? min(min(prm.iterations, MAX_ITERS), first_chunk) : 0;
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
        st.sampled_k = run ? n_first : 0;
        st.sampled_calls = 0; // filled by the sampler thread below
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
    // the sampler needs only M: one thread draws the first chunk while the others sort
    if (tid >= PREP_THREADS - 32) {
      if (tid == PREP_THREADS - 1) {
        RngState s;
        if (per_pair_rng)
            rng_seed(s, seed + (unsigned)(first_pair_index + pair));
        else
            s = *ctx_rng;
        int calls = 0;
        sample_chunk(s, M, ss, 0, n_first, samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE, rng_used_all + (size_t)pair * MAX_ITERS, calls);
        pair_rng[pair] = s; // working copy, continued by k_se3_replay when more chunks are needed
        replay[pair].sampled_calls = calls;
      }
    } else
        cta_std_sort(keys, M, Lpos, Rpos, leaves, qa, qb, s_counters, PREP_THREADS - 32); // :52
    __syncthreads();
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
        if (isnan(from.z) || isnan(to.z))
            wt = 0.0f;
        w[j] = wt;
    }
    __syncthreads();
    // depthCovariance latch (:282-287): the first source depth that reaches errorFunction2
    if (tid == 0 && !latch->latched) {
        for (int j = 0; j < M; j++) {
            const rgbdfe_point3f o = P1[j], t = P2[j];
            if (o.z == 0.0f || t.x == 0.0f)
                continue;
            if (isnan(o.z) || isnan(t.z))
                continue;
            const double z = (double)o.z;
            const double sd = 0.01 * z * z;
            latch->cov = sd * sd;
            __threadfence();
            latch->latched = 1;
            break;
        }
    }
}

// pcl::TransformationFromCorrespondences over the inlier list, one warp.  The 9 covariance
// entries are independent recurrences given (alpha, d1, d2): lane c < 9 owns cov(i = c/3, j = c%3)
// and tracks mean2_i and mean1_j itself, so every value goes through exactly the operations of
// the sequential `add` (same order of points) while the dependent chain per point shrinks to
// one component.  Lanes >= 9 mirror lane % 9.
__device__ __forceinline__ void warp_tfc_fit(const float* __restrict__ P1f, const float* __restrict__ P2f, const float* __restrict__ w,
    const int* __restrict__ idx, int cnt, int lane, float* T)
{
    const int c = lane % 9, ci = c / 3, cj = c % 3;
    float accW = 0.0f, m1 = 0.0f, m2 = 0.0f, cv = 0.0f;
    for (int k = 0; k < cnt; k++) {
        const int jj = idx[k];
        const float wt = w[jj];
        if (wt == 0.0f)
            continue; // NaN depth (SolverSE3.cpp:169-170) or weight == 0 (tfc.add early return)
        const float from = P1f[jj * 3 + cj], to = P2f[jj * 3 + ci];
        accW += wt;
        const float alpha = wt / accW;
        const float d1 = from - m1, d2 = to - m2;
        const float oma = 1.0f - alpha;
        const float ad2 = alpha * d2;
        cv = oma * (cv + ad2 * d1);
        m1 += alpha * d1;
        m2 += alpha * d2;
    }
    Tfc tfc;
    tfc.accW = accW;
#pragma unroll
    for (int e = 0; e < 9; e++)
        tfc.cov[e] = __shfl_sync(0xffffffffu, cv, e);
#pragma unroll
    for (int e = 0; e < 3; e++) {
        tfc.mean1[e] = __shfl_sync(0xffffffffu, m1, e);     // lanes (0, j)
        tfc.mean2[e] = __shfl_sync(0xffffffffu, m2, 3 * e); // lanes (i, 0)
    }
    tfc.transform(T);
}

// dynamic smem: double d2[cap]; int idx[cap]; unsigned char flag[cap]; unsigned cur_mask[mw], ref_mask[mw]
__global__ void __launch_bounds__(HYP_THREADS) k_se3_hypotheses(const rgbdfe_point3f* __restrict__ P1_all, const rgbdfe_point3f* __restrict__ P2_all,
    const float* __restrict__ w_all, int match_cap, const int* __restrict__ nmatch, Se3Params prm, const DepthLatch* __restrict__ latch,
    const int* __restrict__ samples_all, HypOut* __restrict__ hyp_all, unsigned* __restrict__ mask_all, int mask_words,
    const ReplayState* __restrict__ replay, int hyp_begin, int hyp_end)
{
    const int pair = blockIdx.y;
    const ReplayState st = replay[pair];
    const int h = hyp_begin + blockIdx.x;
    if (st.done || !st.run || h >= hyp_end || h >= st.sampled_k)
        return;
    const int M = nmatch[pair];
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sd2 = reinterpret_cast<double*>(smem_raw);
    int* idx = reinterpret_cast<int*>(sd2 + match_cap);
    unsigned* cur_mask = reinterpret_cast<unsigned*>(idx + match_cap);
    unsigned* ref_mask = cur_mask + mask_words;
    unsigned char* flag = reinterpret_cast<unsigned char*>(ref_mask + mask_words);
    __shared__ float sT[16], sRefT[16];
    __shared__ int s_cnt, s_stop, s_refCount;
    __shared__ double s_refErr;

    const rgbdfe_point3f* P1 = P1_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P2 = P2_all + (size_t)pair * match_cap;
    const float* P1f = reinterpret_cast<const float*>(P1);
    const float* P2f = reinterpret_cast<const float*>(P2);
    const float* w = w_all + (size_t)pair * match_cap;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int words = (M + 31) >> 5;
    if (tid == 0) {
        const int* s = samples_all + ((size_t)pair * MAX_ITERS + h) * SE3_MAX_SAMPLE;
        int c = 0;
        for (int j = 0; j < SE3_MAX_SAMPLE; j++)
            if (s[j] >= 0)
                idx[c++] = s[j];
        s_cnt = c;
        s_stop = 0;
        s_refCount = 0;
        s_refErr = 1e6;
        for (int i = 0; i < 16; i++)
            sRefT[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    }
    for (int i = tid; i < words; i += HYP_THREADS)
        ref_mask[i] = 0;
    __syncthreads();
    const double C = latch->cov;
    const double th = (double)(prm.max_mahal * prm.max_mahal);

    for (int refinements = 1; refinements < 20; refinements++) {
        if (warp == 0) { // getTransformFromMatches (:161-179)
            float T[16];
            warp_tfc_fit(P1f, P2f, w, idx, s_cnt, lane, T);
            if (lane < 16)
                sT[lane] = T[lane];
        }
        __syncthreads();
        // computeInliersAndError (:181-214): score all matches
        double Td[16];
#pragma unroll
        for (int i = 0; i < 16; i++)
            Td[i] = (double)sT[i];
        for (int j = tid; j < M; j += HYP_THREADS) {
            const rgbdfe_point3f o = P1[j], t = P2[j];
            bool in = false;
            double md = 0.0;
            if (!(o.z == 0.0f || t.x == 0.0f)) {
                md = mahal2(o, t, Td, C, prm.rcx, prm.rcy);
                in = !(md > th) && (md >= 0.0);
            }
            flag[j] = in;
            sd2[j] = md;
        }
        __syncthreads();
        if (warp == 0) {
            // ordered inlier list + bitmask, then the FP64 error sum in match order
            int base = 0;
            for (int wd = 0; wd < words; wd++) {
                const int j = wd * 32 + lane;
                const bool in = j < M && flag[j];
                const unsigned bal = __ballot_sync(0xffffffffu, in);
                if (in)
                    idx[base + __popc(bal & ((1u << lane) - 1))] = j;
                if (lane == 0)
                    cur_mask[wd] = bal;
                base += __popc(bal);
            }
            __syncwarp();
            if (lane == 0) {
                const int cnt = base;
                double sum = 0.0;
                for (int k = 0; k < cnt; k++)
                    sum += sd2[idx[k]];
                double err = 1e9;
                if (cnt >= 3)
                    err = sqrt(sum / (double)cnt);
                s_cnt = cnt;
                int stop = 0, accept = 0;
                if (cnt < (int)prm.min_inlier_th || err > (double)prm.max_mahal)
                    stop = 1;
                else if (cnt >= s_refCount && err <= s_refErr) {
                    accept = 1;
                    if (cnt == s_refCount)
                        stop = 1;
                    s_refErr = err;
                    s_refCount = cnt;
                } else
                    stop = 1;
                s_stop = stop | (accept << 1);
            }
        }
        __syncthreads();
        const int ctl = s_stop;
        if (ctl & 2) { // vRefinedMatches = vInlierMatches; refinedTransformation = transformation
            for (int i = tid; i < words; i += HYP_THREADS)
                ref_mask[i] = cur_mask[i];
            if (tid < 16)
                sRefT[tid] = sT[tid];
        }
        __syncthreads();
        if (ctl & 1)
            break;
    }
    HypOut* out = hyp_all + (size_t)pair * MAX_ITERS + h;
    if (tid == 0) {
        out->count = s_refCount;
        out->err = s_refErr;
    }
    if (tid < 16)
        out->T[tid] = sRefT[tid];
    unsigned* om = mask_all + ((size_t)pair * MAX_ITERS + h) * mask_words;
    for (int i = tid; i < words; i += HYP_THREADS)
        om[i] = ref_mask[i];
}

// One warp per pair.  dynamic smem: unsigned mask[mask_words] (identity fallback)
__global__ void __launch_bounds__(32) k_se3_replay(const MatchPair* __restrict__ pairs, const rgbdfe_dmatch* __restrict__ sorted_all,
    const rgbdfe_point3f* __restrict__ P1_all, const rgbdfe_point3f* __restrict__ P2_all, int match_cap, Se3Params prm, int update_f2,
    int per_pair_rng, RngState* __restrict__ ctx_rng, RngState* __restrict__ pair_rng, const DepthLatch* __restrict__ latch,
    int* __restrict__ samples_all, int* __restrict__ rng_used_all, const HypOut* __restrict__ hyp_all, const unsigned* __restrict__ mask_all,
    int mask_words, ReplayState* __restrict__ replay, int hyp_end, int next_end, rgbdfe_se3_result* __restrict__ results,
    int* __restrict__ inl_pos_all)
{
    extern __shared__ unsigned s_mask[];
    const int pair = blockIdx.x, lane = threadIdx.x;
    ReplayState st = replay[pair];
    if (st.done == 2)
        return; // already finalised in an earlier stage
    const int M = st.nm;
    const MatchPair P = pairs[pair];
    const rgbdfe_dmatch* m = sorted_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P1 = P1_all + (size_t)pair * match_cap;
    const rgbdfe_point3f* P2 = P2_all + (size_t)pair * match_cap;
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
    const int max_iters = min(prm.iterations, MAX_ITERS);
    if (!st.done) {
        // every lane replays redundantly (uniform control flow)
        bool broke = false;
        const HypOut* hyp = hyp_all + (size_t)pair * MAX_ITERS;
        while (st.n < max_iters && Mok && st.k < hyp_end && st.k < st.sampled_k) {
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
        if (broke || st.n >= max_iters || !Mok)
            st.done = 1;
    }
    if (!st.done) {
        // the loop goes on: draw the samples of the next chunk, continuing the pair's stream
        if (lane == 0) {
            const int k1 = min(next_end, max_iters);
            if (k1 > st.sampled_k) {
                RngState s = pair_rng[pair];
                int calls = st.sampled_calls;
                sample_chunk(s, M, (int)prm.sample_size, st.sampled_k, k1, samples_all + (size_t)pair * MAX_ITERS * SE3_MAX_SAMPLE,
                    rng_used_all + (size_t)pair * MAX_ITERS, calls);
                pair_rng[pair] = s;
                st.sampled_calls = calls;
                st.sampled_k = k1;
            }
            replay[pair] = st;
        }
        return;
    }
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
        const double th = (double)(prm.max_mahal * prm.max_mahal);
        double sum = 0.0;
        int cnt = 0;
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
    // chunk boundaries: a single pair fills the GPU with hypotheses; many pairs start small
    const int b1 = npairs == 1 ? 32 : 8, b2 = npairs == 1 ? 160 : 64;
    const int cap = c->kp_cap;
    const size_t prep_smem = (size_t)cap * (sizeof(SKey) + 3 * sizeof(int)) + 2 * SORT_QCAP * sizeof(int3);
    const size_t hyp_smem = (size_t)cap * (8 + 4 + 1) + (size_t)2 * c->mask_words * 4 + 16;
    static size_t conf_prep = 0, conf_hyp = 0;
    if (prep_smem > conf_prep) {
        CUDA_TRY(c, cudaFuncSetAttribute(k_se3_prepare, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prep_smem));
        conf_prep = prep_smem;
    }
    if (hyp_smem > conf_hyp) {
        CUDA_TRY(c, cudaFuncSetAttribute(k_se3_hypotheses, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hyp_smem));
        conf_hyp = hyp_smem;
    }
    {
        ProfScope prof(c, KID_SE3_PREPARE);
        k_se3_prepare<<<npairs, PREP_THREADS, prep_smem, c->stream>>>(d_pairs, c->d_matches, c->d_sorted, c->d_P1, c->d_P2, c->d_w, cap, c->d_nmatch,
            prm, update_f2, per_pair_rng, seed, first_pair_index, min_matches, c->d_rng, c->d_pair_rng, c->d_latch, c->d_samples, c->d_rng_used,
            c->d_replay, b1);
        c->launches++;
    }
    const int bounds[4] = { 0, b1, b2, iters };
    for (int s = 0; s < 3; s++) {
        const int b = bounds[s] < iters ? bounds[s] : iters;
        const int e = (s == 2 || bounds[s + 1] > iters) ? iters : bounds[s + 1];
        const int next_end = s == 0 ? (b2 < iters ? b2 : iters) : iters;
        if (e > b) {
            ProfScope prof(c, KID_SE3_HYP);
            k_se3_hypotheses<<<dim3(e - b, npairs), HYP_THREADS, hyp_smem, c->stream>>>(c->d_P1, c->d_P2, c->d_w, cap, c->d_nmatch, prm, c->d_latch,
                c->d_samples, c->d_hyp, c->d_hyp_mask, c->mask_words, c->d_replay, b, e);
            c->launches++;
        }
        ProfScope prof2(c, KID_SE3_REPLAY);
        k_se3_replay<<<npairs, 32, (size_t)c->mask_words * 4, c->stream>>>(d_pairs, c->d_sorted, c->d_P1, c->d_P2, cap, prm, update_f2, per_pair_rng,
            c->d_rng, c->d_pair_rng, c->d_latch, c->d_samples, c->d_rng_used, c->d_hyp, c->d_hyp_mask, c->mask_words, c->d_replay, e, next_end,
            d_results, c->d_inl_pos);
        c->launches++;
        if (e >= iters)
            break;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
