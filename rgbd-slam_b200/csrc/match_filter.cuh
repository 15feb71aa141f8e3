// K8 — the filter half of Matcher::match, shared by k_ratio_filter (host-array API) and the
// fused k_se3_prepare of the tracking chain.
#pragma once
#include "common.cuh"

namespace rgbdfe {

// Ratio test + filters + unique-train claim + ordered compaction of Matcher::match
// (Features/Matcher.cpp:115-136) for pair `pair`, run by one CTA of 256 threads.
// s_claim: nt ints of shared memory.  Returns the match count (same value in every thread).
__device__ inline int cta_ratio_filter(const MatchPair& P, int pair, int* s_claim, int4* partial, int q_cap, int splits,
    float ratio, int discard_outliers, rgbdfe_dmatch* matches, int match_cap, int* knn_out)
{
    __shared__ int s_scan[8];
    __shared__ int s_carry;
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
                const int4 v = partial[((size_t)pair * splits + s) * q_cap + q];
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
                int* ko = knn_out + ((size_t)pair * q_cap + q) * 4;
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
            partial[((size_t)pair * splits) * q_cap + q] = make_int4(d0, i0, pass ? 1 : 0, i1);
        }
    }
    __syncthreads();
    // pass 2: ordered compaction (ascending query index)
    for (int q0 = 0; q0 < nq; q0 += 256) {
        const int q = q0 + tid;
        int4 v = make_int4(0, 0, 0, 0);
        bool win = false;
        if (q < nq && enough) {
            v = partial[((size_t)pair * splits) * q_cap + q];
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
                matches[(size_t)pair * match_cap + pos] = m;
            }
        }
        __syncthreads();
        if (tid == 0)
            s_carry = base + total;
        __syncthreads();
    }
    return min(s_carry, match_cap);
}


} // namespace rgbdfe
