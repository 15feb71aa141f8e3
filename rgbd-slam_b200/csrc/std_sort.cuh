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

// Called by the first `nthr` threads of the CTA (a multiple of 32); a: smem keys; scratch: Lpos[n], Rpos[n], leaves[n], a queue of
// SORT_QCAP ranges (f, l, depth) in qa and its publication flags in qb.  Synchronises on named barrier 1.
//
// Ranges are independent once created, so there is no level-by-level barrier: every warp takes a range from a shared queue and
// follows it depth-first -- partition, hand the right part to the queue (another warp picks it up at once), go on with the left
// part -- until what it holds is a leaf.  The critical path is the two or three big partitions at the top; the many small ranges
// below them (17 .. 40 keys: match distances are heavily tied) run concurrently instead of costing a CTA barrier per level.
// Every range > 16 that is ever queued is a distinct node of the recursion tree with >= 17 keys: at most 2 n / 17 < SORT_QCAP
// pushes for n <= 4096, so queue slots are never reused.  counters: [0] head, [1] tail, [2] ranges in flight, [3] leaves.
#define SORT_QCAP 512
#define SORT_BAR() asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory")
__device__ void cta_std_sort(SKey* a, int n, int* Lpos, int* Rpos, int* leaves, int3* qa, int3* qb, int* counters /* [4] */, int nthr, long long* dbg = nullptr)
{
    const int tid = threadIdx.x, lane = tid & 31;
    if (n <= 1)
        return;
    volatile int* vc = counters;
    volatile int* ready = reinterpret_cast<volatile int*>(qb); // ready[slot] = slot + 1 once qa[slot] is written
    for (int i = tid; i < SORT_QCAP; i += nthr)
        ready[i] = 0;
    if (tid == 0) {
        int lg = 0;
        for (int t = n; t > 1; t >>= 1)
            lg++;
        counters[0] = 0, counters[3] = 0;
        if (n > 16) {
            qa[0] = make_int3(0, n, 2 * lg);
            counters[1] = 1, counters[2] = 1;
        } else {
            leaves[0] = 0 | (n << 16);
            counters[1] = 0, counters[2] = 0, counters[3] = 1;
        }
    }
    SORT_BAR();
    if (tid == 0 && n > 16)
        ready[0] = 1;
    if (dbg && tid == 0)
        dbg[0] = dbg_now();
    for (;;) {
        // ---- take a range ----
        int f = 0, l = 0, depth = -1;
        if (lane == 0) {
            for (;;) {
                const int h = vc[0];
                if (h < vc[1]) {
                    if (atomicCAS(&counters[0], h, h + 1) == h) {
                        while (ready[h] != h + 1) { }
                        __threadfence_block();
                        const int3 rg = qa[h];
                        f = rg.x, l = rg.y, depth = rg.z;
                        break;
                    }
                } else if (vc[2] == 0)
                    break; // nothing queued and nothing in flight: the sort is done
            }
        }
        f = __shfl_sync(0xffffffffu, f, 0), l = __shfl_sync(0xffffffffu, l, 0), depth = __shfl_sync(0xffffffffu, depth, 0);
        if (depth < 0)
            break;
        // ---- follow it depth-first ----
        for (;;) {
            if (l - f <= 16) {
                if (lane == 0 && l - f > 1)
                    leaves[atomicAdd(&counters[3], 1)] = f | (l << 16);
                break;
            }
            if (depth == 0) { // heapsort fallback of __introsort_loop; the range is then fully sorted
                if (lane == 0)
                    heap_sort_(a + f, a + l, KeyLess());
                __syncwarp();
                break;
            }
            int cut;
            warp_partition_range(a, f, l, Lpos, Rpos, lane, &cut);
            depth--;
            // right part [cut, l): to the queue when it needs another partition, a leaf otherwise; this warp keeps the left part
            if (lane == 0) {
                if (l - cut > 16) {
                    atomicAdd(&counters[2], 1);
                    const int slot = atomicAdd(&counters[1], 1);
                    if (slot < SORT_QCAP) {
                        qa[slot] = make_int3(cut, l, depth);
                        __threadfence_block();
                        ready[slot] = slot + 1;
                    }
                } else if (l - cut > 1)
                    leaves[atomicAdd(&counters[3], 1)] = cut | (l << 16);
            }
            l = cut;
        }
        if (lane == 0) {
            __threadfence_block();
            atomicSub(&counters[2], 1);
        }
        __syncwarp();
    }
    SORT_BAR();
    if (dbg && tid == 0)
        dbg[2] = dbg_now(), dbg[3] = -counters[3];
    // final insertion sort, leaf by leaf (stable insertion == __final_insertion_sort restricted to the leaf)
    const int nleaves = counters[3];
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
