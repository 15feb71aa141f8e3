// K7 k_knn2_hamming + K8 k_ratio_filter — Matcher::match (Features/Matcher.cpp:106-139):
// cv::BFMatcher(NORM_HAMMING)::knnMatch(k=2), ratio test, unique-train claim,
// outlier / valid-depth filters.  Integer-pipe bound: the query descriptor lives in
// 8 registers of its thread, train descriptors are staged in shared memory and read as
// warp-wide broadcasts (2 x LDS.128 per descriptor per warp).  POPC issues on the XU pipe at a
// quarter of the LOP3 rate, so the 256-bit count goes through carry-save adders first
// (hamming256): 5 POPC + 14 LOP3 per descriptor pair instead of 8 POPC + 8 LOP3.
// Neighbour order is lexicographic (distance, trainIdx) as in OpenCV.
#include "match_filter.cuh"

namespace rgbdfe {

#define KNN_THREADS 128
#define KNN_TILE 128
#ifndef KNN_POPC4
#define KNN_POPC4 1
#endif

// carry-save adder on 32 bit lanes: (a, b, c) -> sum bits and carry bits (one LOP3 each)
__device__ __forceinline__ void csa(unsigned a, unsigned b, unsigned c, unsigned& s, unsigned& cy)
{
    s = a ^ b ^ c;
    cy = (a & b) | (a & c) | (b & c);
}

// popcount of the 256-bit XOR of two descriptors
__device__ __forceinline__ int hamming256(const uint4& qa, const uint4& qb, const uint4& ta, const uint4& tb)
{
    const unsigned x0 = qa.x ^ ta.x, x1 = qa.y ^ ta.y, x2 = qa.z ^ ta.z, x3 = qa.w ^ ta.w;
    const unsigned x4 = qb.x ^ tb.x, x5 = qb.y ^ tb.y, x6 = qb.z ^ tb.z, x7 = qb.w ^ tb.w;
    unsigned s1, c1, s2, c2, s3, c3, t1, f1;
    csa(x0, x1, x2, s1, c1);
    csa(x3, x4, x5, s2, c2);
    csa(s1, s2, x6, s3, c3);
#if KNN_POPC4
    csa(c1, c2, c3, t1, f1);
    return __popc(s3) + __popc(x7) + 2 * __popc(t1) + 4 * __popc(f1);
#else
    return __popc(s3) + __popc(x7) + 2 * (__popc(c1) + __popc(c2) + __popc(c3));
#endif
}

// partial: [pair][split][q] -> int4 (d0, i0, d1, i1)
__global__ void __launch_bounds__(KNN_THREADS) k_knn2_hamming(const MatchPair* __restrict__ pairs, int slot0, int4* __restrict__ partial, int q_cap,
    int splits)
{
    const int pair = slot0 + blockIdx.z;
    const MatchPair P = pairs[pair];
    const int nq = *P.nq, nt = *P.nt;
    const int q = blockIdx.x * KNN_THREADS + threadIdx.x;
    if (blockIdx.x * KNN_THREADS >= nq)
        return;
    __shared__ uint4 s_t[KNN_TILE * 2];
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (q < nq) {
        const uint4* qp = reinterpret_cast<const uint4*>(P.q_desc + (size_t)q * 32);
        qa = qp[0];
        qb = qp[1];
    }
    // train range of this split, in whole tiles
    const int tiles = (nt + KNN_TILE - 1) / KNN_TILE;
    const int tps = (tiles + splits - 1) / splits;
    const int t_begin = min(blockIdx.y * tps * KNN_TILE, nt), t_end = min((blockIdx.y + 1) * tps * KNN_TILE, nt);
    int d0 = 0x7fffffff, d1 = 0x7fffffff, i0 = -1, i1 = -1;
    for (int t0 = t_begin; t0 < t_end; t0 += KNN_TILE) {
        const int cnt = min(KNN_TILE, t_end - t0);
        __syncthreads();
        if ((int)threadIdx.x < cnt) {
            const uint4* tp = reinterpret_cast<const uint4*>(P.t_desc + (size_t)(t0 + threadIdx.x) * 32);
            s_t[threadIdx.x * 2] = tp[0];
            s_t[threadIdx.x * 2 + 1] = tp[1];
        }
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const uint4 ta = s_t[j * 2], tb = s_t[j * 2 + 1];
            const int d = hamming256(qa, qb, ta, tb);
            if (d < d1) { // ascending train index: strict < keeps the lower index on ties
                if (d < d0) {
                    d1 = d0;
                    i1 = i0;
                    d0 = d;
                    i0 = t0 + j;
                } else {
                    d1 = d;
                    i1 = t0 + j;
                }
            }
        }
    }
    if (q < nq)
        partial[((size_t)pair * splits + blockIdx.y) * q_cap + q] = make_int4(d0, i0, d1, i1);
}

// One CTA per pair.  dynamic smem: int claim[t_cap]
__global__ void __launch_bounds__(256) k_ratio_filter(const MatchPair* __restrict__ pairs, int slot0, int4* partial, int q_cap,
    int splits, float ratio, int discard_outliers, rgbdfe_dmatch* __restrict__ matches, int match_cap, int* __restrict__ nmatch,
    int* __restrict__ knn_out)
{
    extern __shared__ int s_claim_dyn[];
    const int pair = slot0 + blockIdx.x;
    const MatchPair P = pairs[pair];
    const int n = cta_ratio_filter(P, pair, s_claim_dyn, partial, q_cap, splits, ratio, discard_outliers, matches, match_cap, knn_out);
    if (threadIdx.x == 0)
        nmatch[pair] = n;
}

int knn_splits_for(const rgbdfe_ctx* c, int npairs) { return npairs >= 64 ? 1 : c->knn_splits; }

// K7 for pair slots [slot0, slot0 + npairs)
int launch_knn(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    dim3 grd((max_nq + KNN_THREADS - 1) / KNN_THREADS, splits, npairs);
    ProfScope prof(c, KID_KNN);
    k_knn2_hamming<<<grd, KNN_THREADS, 0, c->stream>>>(d_pairs, slot0, c->d_partial, c->kp_cap, splits);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

int launch_match(rgbdfe_ctx* c, const MatchPair* d_pairs, int npairs, int max_nq, int max_nt, float ratio, int discard, int* d_knn_out)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    // few pairs: split the train range over several CTAs for latency; many pairs: one split
    const int splits = knn_splits_for(c, npairs);
    int rc = launch_knn(c, d_pairs, 0, npairs, max_nq, splits);
    if (rc != RGBDFE_OK)
        return rc;
    ProfScope prof(c, KID_FILTER);
    size_t smem = (size_t)max_nt * sizeof(int);
    k_ratio_filter<<<npairs, 256, smem, c->stream>>>(d_pairs, 0, c->d_partial, c->kp_cap, splits, ratio, discard, c->d_matches, c->kp_cap, c->d_nmatch,
        d_knn_out);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
