// Parallel, permutation-exact std::sort for the match list (Solver/SolverSE3.cpp:52).
#pragma once
#include "introsort.cuh"

namespace rgbdfe {

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

__device__ void warp_partition_range(SKey* a, int f, int l, int* Lpos, int* Rpos, int lane, int* cut_out, long long* dbg = nullptr)
{
    if (lane != 0)
        dbg = nullptr;
    if (dbg) dbg[0] = dbg_now();
    if (lane == 0) {
        int mid = f + (l - f) / 2;
        move_median_to_first_(a + f, a + f + 1, a + mid, a + l - 1, KeyLess());
    }
    __syncwarp();
    if (dbg) dbg[1] = dbg_now();
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
    if (dbg) dbg[2] = dbg_now();
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
    if (dbg) dbg[3] = dbg_now();
    for (int k = lane; k < K; k += 32) {
        const int x = Lpos[f + k], y = Rpos[f + k];
        const SKey t = a[x];
        a[x] = a[y];
        a[y] = t;
    }
    __syncwarp();
    if (dbg) dbg[4] = dbg_now();
    *cut_out = cut;
}

// All PREP_THREADS threads call this.  a: smem keys; scratch: Lpos[n], Rpos[n], leaves[n],
// lists: 2 x QCAP ranges (f, l, depth packed in int3).
#define SORT_QCAP 512
// Called by the first `nthr` threads of the CTA (a multiple of 32); synchronises on named barrier 1.
#define SORT_BAR() asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory")
__device__ void cta_std_sort(SKey* a, int n, int* Lpos, int* Rpos, int* leaves, int3* qa, int3* qb, int* counters /* [3] */, int nthr, long long* dbg = nullptr)
{
    int dbg_level = 0;
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
        if (dbg && tid == 0 && dbg_level < 40) {
            dbg[2 * dbg_level] = dbg_now();
            dbg[2 * dbg_level + 1] = ncur;
            dbg_level++;
        }
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
            warp_partition_range(a, f, l, Lpos, Rpos, lane, &cut, (dbg && r == 0 && dbg_level < 12) ? dbg + 80 + 8 * dbg_level : nullptr);
            if (dbg && r == 0 && lane == 0 && dbg_level < 12) dbg[80 + 8 * dbg_level + 6] = l - f;
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
        if (dbg && tid == 0 && dbg_level < 12) dbg[80 + 8 * dbg_level + 5] = dbg_now();
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
    if (dbg && tid == 0 && dbg_level < 40) {
        dbg[2 * dbg_level] = dbg_now();
        dbg[2 * dbg_level + 1] = -counters[1];
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

} // namespace rgbdfe
