// cv::KeyPointsFilter::retainBest (OpenCV features2d/src/keypoint.cpp) as cv::ORB uses it (row N3,
// Features/Extractor.cpp:55-59,163-166): std::nth_element + std::partition.  Which keypoints survive is defined by
// the responses; the ORDER they come out in is libstdc++'s introselect permutation (bits/stl_algo.h, GCC 13:
// __introselect with median-of-3 to first, unguarded partition, depth limit 2*floor(log2 n), __heap_select
// fallback, final insertion sort on <= 3 elements) followed by the bidirectional std::partition.  Restated here
// so that host (tests/cpp/test_introselect.cpp checks it against the real library) and device produce that order.
#pragma once
#include "introsort.cuh"

namespace rgbdfe {

template <class T, class C>
__host__ __device__ inline void heap_select_(T* first, T* middle, T* last, C comp)
{
    // std::__heap_select: make_heap(first, middle), then every later element that beats the top replaces it
    const int len = (int)(middle - first);
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            T value = first[parent];
            adjust_heap_(first, parent, len, value, comp);
            if (parent == 0)
                break;
            parent--;
        }
    }
    for (T* i = middle; i < last; ++i)
        if (comp(*i, *first)) {
            T value = *i; // __pop_heap(first, middle, i)
            *i = *first;
            adjust_heap_(first, 0, len, value, comp);
        }
}

// std::nth_element(first, first + nth, last, comp), one thread
template <class T, class C>
__host__ __device__ inline void std_nth_element(T* base, int nth, int n, C comp)
{
    if (n == 0 || nth == n)
        return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1)
        lg++;
    int depth = 2 * lg;
    int f = 0, l = n;
    while (l - f > 3) {
        if (depth == 0) {
            heap_select_(base + f, base + nth + 1, base + l, comp);
            isort_swap(base[f], base[nth]);
            return;
        }
        --depth;
        const int mid = f + (l - f) / 2;
        move_median_to_first_(base + f, base + f + 1, base + mid, base + l - 1, comp);
        const int cut = (int)(unguarded_partition_(base + f + 1, base + l, base + f, comp) - base);
        if (cut <= nth)
            f = cut;
        else
            l = cut;
    }
    insertion_sort_(base + f, base + l, comp);
}

// std::partition(first, last, pred) for bidirectional iterators; returns the index of the partition point
template <class T, class P>
__host__ __device__ inline int std_partition(T* base, int first, int last, P pred)
{
    while (true) {
        while (true) {
            if (first == last)
                return first;
            else if (pred(base[first]))
                ++first;
            else
                break;
        }
        --last;
        while (true) {
            if (first == last)
                return first;
            else if (!pred(base[last]))
                --last;
            else
                break;
        }
        isort_swap(base[first], base[last]);
        ++first;
    }
}

#ifdef __CUDACC__
// Warp-parallel forms that produce the same arrays.  Stopper positions (relative to the range start) are kept as
// 16-bit values: ranges are limited to 65535 elements by the caller.
//
// __unguarded_partition_pivot(first = a + f, last = a + l): L-stoppers (!comp(a[x], pivot)) in ascending position and
// R-stoppers (!comp(pivot, a[x])) in descending position are compacted with ballots; the k-th swap of the sequential
// loop exchanges the k-th of each list while L[k] < R[k]; the cut is min(L[K], R[K-1]) (same construction as
// std_sort.cuh, which tests/cpp/test_introsort.cpp checks against the library on the host).
template <class T, class C>
__device__ int warp_partition_pivot(T* a, int f, int l, unsigned short* Lpos, unsigned short* Rpos, int lane, C comp)
{
    if (lane == 0) {
        const int mid = f + (l - f) / 2;
        move_median_to_first_(a + f, a + f + 1, a + mid, a + l - 1, comp);
    }
    __syncwarp();
    const T p = a[f];
    int nL = 0, nR = 0;
    const int first = f + 1;
    const unsigned lt = (1u << lane) - 1;
    for (int x0 = first; x0 < l; x0 += 32) {
        const int x = x0 + lane;
        const bool isL = x < l && !comp(a[x], p);
        const unsigned bal = __ballot_sync(0xffffffffu, isL);
        if (isL)
            Lpos[nL + __popc(bal & lt)] = (unsigned short)(x - f);
        nL += __popc(bal);
    }
    for (int x0 = l - 1; x0 >= first; x0 -= 32) {
        const int x = x0 - lane;
        const bool isR = x >= first && !comp(p, a[x]);
        const unsigned bal = __ballot_sync(0xffffffffu, isR);
        if (isR)
            Rpos[nR + __popc(bal & lt)] = (unsigned short)(x - f);
        nR += __popc(bal);
    }
    __syncwarp();
    const int nmin = min(nL, nR);
    int K = nmin;
    for (int k0 = 0; k0 < nmin; k0 += 32) {
        const int k = k0 + lane;
        const bool stop = k < nmin && !(Lpos[k] < Rpos[k]);
        const unsigned bal = __ballot_sync(0xffffffffu, stop);
        if (bal) {
            K = k0 + __ffs(bal) - 1;
            break;
        }
    }
    const int Lk = K < nL ? f + Lpos[K] : l;
    const int Rprev = K > 0 ? f + Rpos[K - 1] : l;
    const int cut = min(Lk, Rprev);
    __syncwarp();
    for (int k = lane; k < K; k += 32) {
        const int x = f + Lpos[k], y = f + Rpos[k];
        const T t = a[x];
        a[x] = a[y];
        a[y] = t;
    }
    __syncwarp();
    return cut;
}

// std::partition(base + first, base + last, pred) by one warp: the k-th swap exchanges the k-th element failing pred
// (ascending) with the k-th element passing it (descending) while the former lies before the latter.
template <class T, class P>
__device__ int warp_partition(T* base, int first, int last, unsigned short* Lpos, unsigned short* Rpos, int lane, P pred)
{
    int nL = 0, nR = 0;
    const unsigned lt = (1u << lane) - 1;
    for (int x0 = first; x0 < last; x0 += 32) {
        const int x = x0 + lane;
        const bool isL = x < last && !pred(base[x]);
        const unsigned bal = __ballot_sync(0xffffffffu, isL);
        if (isL)
            Lpos[nL + __popc(bal & lt)] = (unsigned short)(x - first);
        nL += __popc(bal);
    }
    for (int x0 = last - 1; x0 >= first; x0 -= 32) {
        const int x = x0 - lane;
        const bool isR = x >= first && pred(base[x]);
        const unsigned bal = __ballot_sync(0xffffffffu, isR);
        if (isR)
            Rpos[nR + __popc(bal & lt)] = (unsigned short)(x - first);
        nR += __popc(bal);
    }
    __syncwarp();
    const int nmin = min(nL, nR);
    int K = nmin;
    for (int k0 = 0; k0 < nmin; k0 += 32) {
        const int k = k0 + lane;
        const bool stop = k < nmin && !(Lpos[k] < Rpos[k]);
        const unsigned bal = __ballot_sync(0xffffffffu, stop);
        if (bal) {
            K = k0 + __ffs(bal) - 1;
            break;
        }
    }
    __syncwarp();
    for (int k = lane; k < K; k += 32) {
        const int x = first + Lpos[k], y = first + Rpos[k];
        const T t = base[x];
        base[x] = base[y];
        base[y] = t;
    }
    __syncwarp();
    return first + nR; // every element passing pred ends up in front
}
// ---- CTA-wide forms: every thread of the CTA calls them (barriers inside); s_w = int[68] of shared memory. ----
// The warps scan disjoint chunks of the range: stopper counts per chunk, a prefix over the chunks, then the positions
// are written in ascending order for both lists (the R list is read back to front: R_desc[k] = Rasc[nR - 1 - k]).
#define CTA_PARTITION_MIN 1024 // shorter ranges are partitioned by warp 0 alone
template <class FL, class FR>
__device__ __forceinline__ void cta_stopper_lists(int first, int l, int f, unsigned short* Lpos, unsigned short* Rasc, int* s_w, FL isLf, FR isRf,
    int* nL_out, int* nR_out)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int m = l - first;
    const int cs = (((m + nwarps - 1) / nwarps) + 31) & ~31;
    const int cb = first + warp * cs, ce = min(cb + cs, l);
    int cl = 0, cr = 0;
    for (int x0 = cb; x0 < ce; x0 += 32) {
        const int x = x0 + lane;
        cl += __popc(__ballot_sync(0xffffffffu, x < ce && isLf(x)));
        cr += __popc(__ballot_sync(0xffffffffu, x < ce && isRf(x)));
    }
    if (lane == 0) {
        s_w[warp] = cl;
        s_w[32 + warp] = cr;
    }
    if (tid == 0)
        s_w[64] = 0x7fffffff;
    __syncthreads();
    int baseL = 0, baseR = 0, nL = 0, nR = 0;
    for (int w = 0; w < nwarps; w++) {
        const int a = s_w[w], b = s_w[32 + w];
        if (w < warp) {
            baseL += a;
            baseR += b;
        }
        nL += a;
        nR += b;
    }
    const unsigned lt = (1u << lane) - 1;
    for (int x0 = cb; x0 < ce; x0 += 32) {
        const int x = x0 + lane;
        const bool isL = x < ce && isLf(x), isR = x < ce && isRf(x);
        const unsigned bl = __ballot_sync(0xffffffffu, isL), br = __ballot_sync(0xffffffffu, isR);
        if (isL)
            Lpos[baseL + __popc(bl & lt)] = (unsigned short)(x - f);
        if (isR)
            Rasc[baseR + __popc(br & lt)] = (unsigned short)(x - f);
        baseL += __popc(bl);
        baseR += __popc(br);
    }
    __syncthreads();
    *nL_out = nL;
    *nR_out = nR;
}

// number of leading k with Lpos[k] < R_desc[k] (the swaps the sequential loop performs), by all threads
__device__ __forceinline__ int cta_swap_count(const unsigned short* Lpos, const unsigned short* Rasc, int nL, int nR, int* s_w)
{
    const int nmin = min(nL, nR);
    for (int k = threadIdx.x; k < nmin; k += blockDim.x)
        if (!(Lpos[k] < Rasc[nR - 1 - k]))
            atomicMin(&s_w[64], k);
    __syncthreads();
    return min(s_w[64], nmin);
}

template <class T, class C>
__device__ int cta_partition_pivot(T* a, int f, int l, unsigned short* Lpos, unsigned short* Rasc, int* s_w, C comp)
{
    if (threadIdx.x == 0) {
        const int mid = f + (l - f) / 2;
        move_median_to_first_(a + f, a + f + 1, a + mid, a + l - 1, comp);
    }
    __syncthreads();
    const T p = a[f];
    int nL, nR;
    cta_stopper_lists(f + 1, l, f, Lpos, Rasc, s_w, [&](int x) { return !comp(a[x], p); }, [&](int x) { return !comp(p, a[x]); }, &nL, &nR);
    const int K = cta_swap_count(Lpos, Rasc, nL, nR, s_w);
    const int Lk = K < nL ? f + Lpos[K] : l;
    const int Rprev = K > 0 ? f + Rasc[nR - K] : l;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const int x = f + Lpos[k], y = f + Rasc[nR - 1 - k];
        const T t = a[x];
        a[x] = a[y];
        a[y] = t;
    }
    __syncthreads();
    return min(Lk, Rprev);
}

template <class T, class C>
__device__ void cta_nth_element(T* base, int nth, int n, unsigned short* Lpos, unsigned short* Rpos, int* s_w, C comp)
{
    if (n == 0 || nth == n)
        return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1)
        lg++;
    int depth = 2 * lg, f = 0, l = n, it = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    while (l - f > 3) {
        if (depth == 0) {
            if (threadIdx.x == 0) {
                heap_select_(base + f, base + nth + 1, base + l, comp);
                isort_swap(base[f], base[nth]);
            }
            __syncthreads();
            return;
        }
        --depth;
        int cut;
        if (l - f >= CTA_PARTITION_MIN)
            cut = cta_partition_pivot(base, f, l, Lpos, Rpos, s_w, comp);
        else {
            if (warp == 0) {
                cut = warp_partition_pivot(base, f, l, Lpos, Rpos, lane, comp);
                if (lane == 0)
                    s_w[65 + (it & 1)] = cut;
            }
            __syncthreads();
            cut = s_w[65 + (it & 1)];
            it++;
        }
        if (cut <= nth)
            f = cut;
        else
            l = cut;
    }
    if (threadIdx.x == 0)
        insertion_sort_(base + f, base + l, comp);
    __syncthreads();
}

template <class T, class P>
__device__ int cta_partition(T* base, int first, int last, unsigned short* Lpos, unsigned short* Rasc, int* s_w, P pred)
{
    if (last - first < CTA_PARTITION_MIN) {
        if ((threadIdx.x >> 5) == 0) {
            const int e = warp_partition(base, first, last, Lpos, Rasc, threadIdx.x & 31, pred);
            if (threadIdx.x == 0)
                s_w[67] = e;
        }
        __syncthreads();
        return s_w[67];
    }
    int nL, nR;
    cta_stopper_lists(first, last, first, Lpos, Rasc, s_w, [&](int x) { return !pred(base[x]); }, [&](int x) { return pred(base[x]); }, &nL, &nR);
    const int K = cta_swap_count(Lpos, Rasc, nL, nR, s_w);
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const int x = first + Lpos[k], y = first + Rasc[nR - 1 - k];
        const T t = base[x];
        base[x] = base[y];
        base[y] = t;
    }
    __syncthreads();
    return first + nR;
}
#endif


} // namespace rgbdfe
