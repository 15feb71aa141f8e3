// K7 k_knn2_hamming + K8 k_ratio_filter — Matcher::match (Features/Matcher.cpp:106-139):
// cv::BFMatcher(NORM_HAMMING)::knnMatch(k=2), ratio test, unique-train claim,
// outlier / valid-depth filters.  Integer-pipe (POPC) bound: the query descriptor lives in
// 8 registers of its thread, train descriptors are staged in shared memory and read as
// warp-wide broadcasts (2 x LDS.128 per descriptor per warp), 8 XOR + 8 POPC per pair.
// Neighbour order is lexicographic (distance, trainIdx) as in OpenCV.
#include "common.cuh"

namespace rgbdfe {

#define KNN_THREADS 128
#define KNN_TILE 128

// partial: [pair][split][q] -> int4 (d0, i0, d1, i1)
__global__ void __launch_bounds__(KNN_THREADS) k_knn2_hamming(const MatchPair* __restrict__ pairs, int4* __restrict__ partial, int q_cap, int splits)
{
    const MatchPair P = pairs[blockIdx.z];
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
            int d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) + __popc(qb.x ^ tb.x)
                + __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
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
        partial[((size_t)blockIdx.z * splits + blockIdx.y) * q_cap + q] = make_int4(d0, i0, d1, i1);
}

// One CTA per pair.  dynamic smem: int claim[t_cap]
__global__ void __launch_bounds__(256) k_ratio_filter(const MatchPair* __restrict__ pairs, const int4* __restrict__ partial, int q_cap, int splits,
    float ratio, int discard_outliers, rgbdfe_dmatch* __restrict__ matches, int match_cap, int* __restrict__ nmatch, int* __restrict__ knn_out)
{
    extern __shared__ int s_claim[];
    __shared__ int s_scan[256];
    __shared__ int s_carry;
    const MatchPair P = pairs[blockIdx.x];
    const int nq = *P.nq, nt = *P.nt;
    const int tid = threadIdx.x;
    for (int i = tid; i < nt; i += 256)
        s_claim[i] = 0x7fffffff;
    if (tid == 0)
        s_carry = 0;
    __syncthreads();
    // Guard: with < 2 train descriptors the reference reads matchesKnn[i][1] out of range (Matcher.cpp:117); we return no match.
    const bool enough = nt >= 2;
    // pass 1: merge partials, test, claim
    for (int q0 = 0; q0 < nq; q0 += 256) {
        const int q = q0 + tid;
        if (q < nq && enough) {
            int d0 = 0x7fffffff, d1 = 0x7fffffff, i0 = -1, i1 = -1;
            for (int s = 0; s < splits; s++) { // splits cover ascending train ranges
                const int4 v = partial[((size_t)blockIdx.x * splits + s) * q_cap + q];
                const int cd[2] = { v.x, v.z }, ci[2] = { v.y, v.w };
                for (int k = 0; k < 2; k++) {
                    if (ci[k] < 0)
                        continue;
                    if (cd[k] < d1) {
                        if (cd[k] < d0) {
                            d1 = d0, i1 = i0, d0 = cd[k], i0 = ci[k];
                        } else {
                            d1 = cd[k], i1 = ci[k];
                        }
                    }
                }
            }
            if (knn_out) {
                int* ko = knn_out + ((size_t)blockIdx.x * q_cap + q) * 4;
                ko[0] = i0, ko[1] = i1, ko[2] = d0, ko[3] = d1;
            }
            bool pass = (float)d0 < __fmul_rn(ratio, (float)d1);
            if (pass && discard_outliers && P.q_outlier && P.q_outlier[q])
                pass = false;
            if (pass && !(P.q_xyz[q].z > 0.f && P.t_xyz[i0].z > 0.f))
                pass = false;
            if (pass)
                atomicMin(&s_claim[i0], q);
            // stash for pass 2 in the partial buffer of split 0 (x = d0, y = i0, z = pass)
            const_cast<int4*>(partial)[((size_t)blockIdx.x * splits) * q_cap + q] = make_int4(d0, i0, pass ? 1 : 0, i1);
        }
    }
    __syncthreads();
    // pass 2: ordered compaction (ascending query index)
    for (int q0 = 0; q0 < nq; q0 += 256) {
        const int q = q0 + tid;
        int4 v = make_int4(0, 0, 0, 0);
        bool win = false;
        if (q < nq && enough) {
            v = partial[((size_t)blockIdx.x * splits) * q_cap + q];
            win = v.z && s_claim[v.y] == q;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, win);
        if ((tid & 31) == 0)
            s_scan[tid >> 5] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int wv = 0; wv < 8; wv++) {
            const int cw = s_scan[wv];
            before += wv < (tid >> 5) ? cw : 0;
            total += cw;
        }
        const int base = s_carry;
        if (win) {
            const int pos = base + before + __popc(bal & ((1u << (tid & 31)) - 1));
            if (pos < match_cap) {
                rgbdfe_dmatch m;
                m.queryIdx = q;
                m.trainIdx = v.y;
                m.imgIdx = 0;
                m.distance = (float)v.x;
                matches[(size_t)blockIdx.x * match_cap + pos] = m;
            }
        }
        __syncthreads();
        if (tid == 0)
            s_carry = base + total;
        __syncthreads();
    }
    if (tid == 0)
        nmatch[blockIdx.x] = min(s_carry, match_cap);
}

int launch_match(rgbdfe_ctx* c, const MatchPair* d_pairs, int npairs, int max_nq, int max_nt, float ratio, int discard, int* d_knn_out)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    // few pairs: split the train range over several CTAs for latency; many pairs: one split
    const int splits = npairs >= 64 ? 1 : c->knn_splits;
    dim3 grd((max_nq + KNN_THREADS - 1) / KNN_THREADS, splits, npairs);
    {
        ProfScope prof(c, KID_KNN);
        k_knn2_hamming<<<grd, KNN_THREADS, 0, c->stream>>>(d_pairs, c->d_partial, c->kp_cap, splits);
        c->launches++;
    }
    ProfScope prof(c, KID_FILTER);
    size_t smem = (size_t)max_nt * sizeof(int);
    k_ratio_filter<<<npairs, 256, smem, c->stream>>>(d_pairs, c->d_partial, c->kp_cap, splits, ratio, discard, c->d_matches, c->kp_cap, c->d_nmatch,
        d_knn_out);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
