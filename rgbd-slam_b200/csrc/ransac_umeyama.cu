// K12 — class Ransac (Solver/Ransac.cpp:46-181,183-280,429-454): Umeyama RANSAC with the
// Euclidean / adaptive inlier test and the adaptive iteration count.
//
// Sequential in the reference: the rand() stream and `iterationCount`, which shrinks whenever a
// better model is saved.  Everything else is per hypothesis.  So:
//   k_ume_prepare     1 CTA: depth gate z in [0.1, 6] on both points (:72-83) as an ordered
//                     compaction, then one thread draws the samples of all 487 initial
//                     iterations (getRandomMatches :183-208 depends only on the match count).
//   k_ume_hypotheses  1 CTA / iteration: Eigen::umeyama (no scaling) on the usedPairs sampled
//                     correspondences (:210-245), Euclidean test of every match (:247-280),
//                     inlier count + ordered bitmask.
//   k_ume_finish      1 CTA: replays saveBetterModel (:429-447) in iteration order with the
//                     iteration-count table, refits on the best inlier set, re-tests (:159-165),
//                     applies the ratio gate (:168-171) and T.isIdentity() (:55-58), and advances
//                     the context rand() stream by exactly the calls the executed iterations made.
// computeRANSACIteration (:449-454) is double log/pow of the inlier ratio; the ratio can only be
// count / M, so the host evaluates it with the C library for every count and uploads the table
// (bit-equal to the reference's libm, which a device log/pow is not guaranteed to be).
// FP contract as in se3_math.cuh (FP32, no FMA); sums over correspondences run in list order.
#include "rng.cuh"
#include <algorithm>
#include "se3_math.cuh"

namespace rgbdfe {

#define UME_THREADS 128
#define UME_MAX_PAIRS 8
#define UME_RNG_BLOCK 1024

// Eigen::umeyama(src, dst, with_scaling = false) over `n` correspondences given by idx (or
// 0..n-1 when idx == nullptr); src = current-frame points (P2), dst = previous-frame points (P1).
// One warp: lanes 0..2 own the mean components, lanes 0..8 the covariance entries, so every sum
// still runs over the correspondences in list order.
__device__ void warp_umeyama(const float* __restrict__ srcf, const float* __restrict__ dstf, const int* __restrict__ idx, int n, int lane, float* T)
{
    const float one_over_n = 1.0f / (float)n;
    const int c = lane % 9, ci = c / 3, cj = c % 3;
    float smj = 0.0f, dmi = 0.0f; // src mean component cj, dst mean component ci
    for (int k = 0; k < n; k++) {
        const int j = idx ? idx[k] : k;
        smj += srcf[j * 3 + cj];
        dmi += dstf[j * 3 + ci];
    }
    smj *= one_over_n;
    dmi *= one_over_n;
    float sg = 0.0f;
    for (int k = 0; k < n; k++) {
        const int j = idx ? idx[k] : k;
        const float sd = srcf[j * 3 + cj] - smj;
        const float dd = one_over_n * (dstf[j * 3 + ci] - dmi);
        sg += dd * sd;
    }
    if (n == 0)
        sg = nanf("");
    float sigma[9], sm[3], dm[3];
#pragma unroll
    for (int e = 0; e < 9; e++)
        sigma[e] = __shfl_sync(0xffffffffu, sg, e);
#pragma unroll
    for (int e = 0; e < 3; e++) {
        sm[e] = __shfl_sync(0xffffffffu, smj, e);     // lanes (0, j)
        dm[e] = __shfl_sync(0xffffffffu, dmi, 3 * e); // lanes (i, 0)
    }
    float U[9], S[3], V[9];
    jacobi_svd3(sigma, U, S, V);
    const float s2 = (det3(U) * det3(V) < 0.0f) ? -1.0f : 1.0f;
    float R[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            R[i * 3 + j] = (U[i * 3 + 0] * V[j * 3 + 0] + U[i * 3 + 1] * V[j * 3 + 1]) + (U[i * 3 + 2] * s2) * V[j * 3 + 2];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const float rm = (R[i * 3 + 0] * sm[0] + R[i * 3 + 1] * sm[1]) + R[i * 3 + 2] * sm[2];
        T[i * 4 + 0] = R[i * 3 + 0];
        T[i * 4 + 1] = R[i * 3 + 1];
        T[i * 4 + 2] = R[i * 3 + 2];
        T[i * 4 + 3] = dm[i] - rm;
    }
    T[12] = T[13] = T[14] = 0.0f;
    T[15] = 1.0f;
}

// computeMatchInlierRatioEuclidean's per-match test (:261-274)
__device__ __forceinline__ bool ume_inlier(const rgbdfe_point3f& p, const rgbdfe_point3f& c, const float* T, double th0, int adaptive)
{
    float e[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const float rc = T[i * 4 + 0] * c.x + (T[i * 4 + 1] * c.y + T[i * 4 + 2] * c.z);
        e[i] = rc + T[i * 4 + 3];
    }
    double threshold = th0;
    if (adaptive)
        threshold *= (double)p.z;
    const float dx = e[0] - p.x, dy = e[1] - p.y, dz = e[2] - p.z;
    const float nrm = sqrtf(dx * dx + (dy * dy + dz * dz));
    return (double)nrm < threshold;
}

__global__ void __launch_bounds__(256) k_ume_prepare(const rgbdfe_dmatch* __restrict__ matches, int nm, const rgbdfe_point3f* __restrict__ xyz1,
    const rgbdfe_point3f* __restrict__ xyz2, rgbdfe_dmatch* __restrict__ kept, rgbdfe_point3f* __restrict__ P1, rgbdfe_point3f* __restrict__ P2,
    int used_pairs, int iters0, const RngState* __restrict__ ctx_rng, int* __restrict__ samples, int* __restrict__ rng_used,
    UmeState* __restrict__ state)
{
    __shared__ int s_warp[8];
    __shared__ int s_base;
    __shared__ unsigned s_lin[RNG_HIST + UME_RNG_BLOCK];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0)
        s_base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < nm; i0 += 256) {
        const int i = i0 + tid;
        bool keep = false;
        rgbdfe_dmatch m;
        rgbdfe_point3f a, b;
        if (i < nm) {
            m = matches[i];
            a = xyz1[m.queryIdx];
            b = xyz2[m.trainIdx];
            keep = !(a.z < 0.1f || a.z > 6.0f || b.z < 0.1f || b.z > 6.0f);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0)
            s_warp[warp] = __popc(bal);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < warp; w++)
            off += s_warp[w];
        if (keep) {
            const int pos = off + __popc(bal & ((1u << lane) - 1));
            kept[pos] = m;
            P1[pos] = a;
            P2[pos] = b;
        }
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int w = 0; w < 8; w++)
                t += s_warp[w];
            s_base += t;
        }
        __syncthreads();
    }
    if (tid != 0)
        return;
    const int M = s_base;
    state->M = M;
    state->total_calls = 0;
    if (M < used_pairs)
        return; // the host refuses this case (the reference would never leave getRandomMatches)
    // getRandomMatches (:183-208): one rand() per draw, already chosen indices are redrawn
    rng_window_load(ctx_rng, s_lin);
    int calls = 0, pos = UME_RNG_BLOCK;
    for (int k = 0; k < iters0; k++) {
        int ids[UME_MAX_PAIRS];
        int cnt = 0;
        while (cnt < used_pairs) {
            if (pos == UME_RNG_BLOCK) { // refill: slide the window over the consumed block, extend it by another one
                if (calls > 0)
                    for (int t = 0; t < RNG_HIST; t++)
                        s_lin[t] = s_lin[UME_RNG_BLOCK + t];
                rng_extend(s_lin, UME_RNG_BLOCK);
                pos = 0;
            }
            const int id = rng_random_int0(s_lin[RNG_HIST + pos], M);
            pos++;
            calls++;
            bool fresh = true;
            for (int j = 0; j < cnt; j++)
                fresh &= ids[j] != id;
            if (fresh)
                ids[cnt++] = id;
        }
        for (int j = 0; j < UME_MAX_PAIRS; j++)
            samples[k * UME_MAX_PAIRS + j] = j < cnt ? ids[j] : -1;
        rng_used[k] = calls;
    }
    state->total_calls = calls;
}

__global__ void __launch_bounds__(UME_THREADS) k_ume_hypotheses(const rgbdfe_point3f* __restrict__ P1, const rgbdfe_point3f* __restrict__ P2,
    const UmeState* __restrict__ state, const int* __restrict__ samples, int used_pairs, double th0, int adaptive, HypOut* __restrict__ hyp,
    unsigned* __restrict__ masks, int mask_words)
{
    __shared__ float sT[16];
    __shared__ int s_valid, s_count;
    __shared__ int s_idx[UME_MAX_PAIRS];
    const int h = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int M = state->M;
    if (M < used_pairs)
        return;
    if (tid == 0)
        s_count = 0;
    if (tid < UME_MAX_PAIRS)
        s_idx[tid] = samples[h * UME_MAX_PAIRS + tid];
    __syncthreads();
    if (tid < 32) {
        float T[16];
        warp_umeyama(reinterpret_cast<const float*>(P2), reinterpret_cast<const float*>(P1), s_idx, used_pairs, lane, T);
        if (lane < 16)
            sT[lane] = T[lane];
        if (lane == 0)
            s_valid = isnan(T[0]) ? 0 : 1; // computeTransformationModel's NaN test (:240-243)
    }
    __syncthreads();
    HypOut* out = hyp + h;
    if (!s_valid) {
        if (tid == 0) {
            out->count = -1;
            out->err = 0.0;
        }
        return;
    }
    float T[16];
#pragma unroll
    for (int i = 0; i < 16; i++)
        T[i] = sT[i];
    const int words = (M + 31) >> 5;
    int local = 0;
    for (int j0 = 0; j0 < words * 32; j0 += UME_THREADS) {
        const int j = j0 + tid;
        const bool in = j < M && ume_inlier(P1[j], P2[j], T, th0, adaptive);
        const unsigned bal = __ballot_sync(0xffffffffu, in);
        if (lane == 0 && (j >> 5) < words) {
            masks[(size_t)h * mask_words + (j >> 5)] = bal;
            local += __popc(bal);
        }
    }
    if (lane == 0)
        atomicAdd(&s_count, local);
    __syncthreads();
    if (tid == 0) {
        out->count = s_count;
        out->err = 0.0;
    }
    if (tid < 16)
        out->T[tid] = T[tid];
}

// dynamic smem: int idx[cap]
__global__ void __launch_bounds__(256) k_ume_finish(const rgbdfe_point3f* __restrict__ P1, const rgbdfe_point3f* __restrict__ P2,
    const UmeState* __restrict__ state, const HypOut* __restrict__ hyp, const unsigned* __restrict__ masks, int mask_words,
    const int* __restrict__ iters_tab, int iters_min, int iters0, double th0, int adaptive, double min_ratio, const int* __restrict__ rng_used,
    RngState* __restrict__ ctx_rng, rgbdfe_umeyama_result* __restrict__ res, int* __restrict__ inl_pos)
{
    extern __shared__ int s_idx[];
    __shared__ int s_best, s_K, s_n, s_warp[8], s_base;
    __shared__ float sT[16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int M = state->M;
    if (tid == 0) {
        // sequential control of the loop at :91-152 with saveBetterModel's iteration-count update
        int iterationCount = iters0, best = -1, bestCount = 0, i = 0;
        for (; i < iterationCount; i++) {
            const int cnt = hyp[i].count;
            if (cnt > bestCount) { // ratio = cnt / M with a common M: "ratio > bestRatio" == "cnt > bestCount"
                best = i;
                bestCount = cnt;
                const int t = iters_tab[cnt];
                iterationCount = iters_min < t ? iters_min : t;
            }
        }
        s_best = best;
        s_K = i;
        s_base = 0;
    }
    __syncthreads();
    const int best = s_best, K = s_K;
    // bestInlierMatches (ordered) from the winning hypothesis' mask
    int n = 0;
    if (best >= 0) {
        const int words = (M + 31) >> 5;
        const unsigned* mk = masks + (size_t)best * mask_words;
        for (int w0 = 0; w0 < words; w0 += 8) {
            const int w = w0 + warp;
            const unsigned bits = w < words ? mk[w] : 0u;
            if (lane == 0)
                s_warp[warp] = __popc(bits);
            __syncthreads();
            int off = s_base;
            for (int x = 0; x < warp; x++)
                off += s_warp[x];
            if (bits & (1u << lane))
                s_idx[off + __popc(bits & ((1u << lane) - 1))] = w * 32 + lane;
            __syncthreads();
            if (tid == 0) {
                int t = 0;
                for (int x = 0; x < 8; x++)
                    t += s_warp[x];
                s_base += t;
            }
            __syncthreads();
        }
        n = s_base;
    }
    // final refit on the best inliers (:154-156); an empty list gives NaN -> identity
    if (tid < 32) {
        float T[16];
        warp_umeyama(reinterpret_cast<const float*>(P2), reinterpret_cast<const float*>(P1), s_idx, n, lane, T);
        if (isnan(T[0])) {
#pragma unroll
            for (int i = 0; i < 16; i++)
                T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        }
        if (lane < 16)
            sT[lane] = T[lane];
    }
    __syncthreads();
    // re-test the inliers under the refitted model (:157-165); one warp keeps the order
    if (tid < 32) {
        float T[16];
#pragma unroll
        for (int i = 0; i < 16; i++)
            T[i] = sT[i];
        int kept = 0;
        for (int k0 = 0; k0 < n; k0 += 32) {
            const int k = k0 + lane;
            int j = -1;
            bool in = false;
            if (k < n) {
                j = s_idx[k];
                in = ume_inlier(P1[j], P2[j], T, th0, adaptive);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, in);
            if (in)
                inl_pos[kept + __popc(bal & ((1u << lane) - 1))] = j;
            kept += __popc(bal);
        }
        if (lane == 0)
            s_n = kept;
    }
    __syncthreads();
    if (tid == 0) {
        int n_inl = s_n;
        float T[16];
        for (int i = 0; i < 16; i++)
            T[i] = sT[i];
        const float ratio = best >= 0 ? (float)hyp[best].count / (float)M : 0.0f;
        if ((double)ratio < min_ratio) { // :168-171
            for (int i = 0; i < 16; i++)
                T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
            n_inl = 0;
        }
        bool ident = true; // Eigen isIdentity(prec = 1e-5f)
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) {
                const float x = T[i * 4 + j];
                if (i == j) {
                    const float ax = fabsf(x);
                    if (!(fabsf(x - 1.0f) <= 1e-5f * (ax < 1.0f ? ax : 1.0f)))
                        ident = false;
                } else if (!(fabsf(x) <= 1e-5f))
                    ident = false;
            }
        res->ok = ident ? 0 : 1;
        for (int i = 0; i < 16; i++)
            res->T[i] = T[i];
        res->n_inliers = n_inl;
        res->iters_executed = K;
        // the executed iterations consumed rng_used[K-1] rand() calls of the process stream
        const int calls = K > 0 ? rng_used[K - 1] : 0;
        rng_advance(reinterpret_cast<unsigned*>(s_idx), ctx_rng, calls, UME_RNG_BLOCK); // s_idx is free by now
    }
}

int launch_ransac_umeyama(rgbdfe_ctx* c, int nm, const rgbdfe_point3f* d_xyz1, const rgbdfe_point3f* d_xyz2, int used_pairs, int iters0,
    const int* d_iters_tab, int iters_min, double th0, int adaptive, double min_ratio, UmeState* d_state, rgbdfe_umeyama_result* d_res)
{
    cudaStream_t s = c->stream;
    k_ume_prepare<<<1, 256, 0, s>>>(c->d_matches, nm, d_xyz1, d_xyz2, c->d_sorted, c->d_P1, c->d_P2, used_pairs, iters0, c->d_rng, c->d_samples,
        c->d_rng_used, d_state);
    k_ume_hypotheses<<<iters0, UME_THREADS, 0, s>>>(c->d_P1, c->d_P2, d_state, c->d_samples, used_pairs, th0, adaptive, c->d_hyp, c->d_hyp_mask,
        c->mask_words);
    k_ume_finish<<<1, 256, std::max((size_t)c->kp_cap, (size_t)(RNG_HIST + UME_RNG_BLOCK)) * sizeof(int), s>>>(c->d_P1, c->d_P2, d_state, c->d_hyp, c->d_hyp_mask, c->mask_words, d_iters_tab,
        iters_min, iters0, th0, adaptive, min_ratio, c->d_rng_used, c->d_rng, d_res, c->d_inl_pos);
    c->launches += 3;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
