// std::sort as implemented by libstdc++ (bits/stl_algo.h, GCC 13: introsort with
// median-of-3 to first, unguarded partition, threshold 16, depth limit 2*floor(log2 n),
// heapsort fallback, final insertion sort), restated so that one device thread
// produces the SAME permutation as the reference's host-side
// `sort(vUsedMatches.begin(), vUsedMatches.end())` (Solver/SolverSE3.cpp:52), whose
// comparator (cv::DMatch::operator<, distance only) leaves ties in an
// algorithm-dependent order that the biased sampler then indexes.
#pragma once
#include "../../include/rgbdfe.h"

#ifndef __CUDACC__
#define __host__
#define __device__
#endif

namespace rgbdfe {

struct DmLess {
    __host__ __device__ bool operator()(const rgbdfe_dmatch& a, const rgbdfe_dmatch& b) const { return a.distance < b.distance; }
};

template <class T>
__host__ __device__ inline void isort_swap(T& a, T& b)
{
    T t = a;
    a = b;
    b = t;
}

template <class T, class C>
__host__ __device__ inline void push_heap_(T* first, int hole, int top, T value, C comp)
{
    int parent = (hole - 1) / 2;
    while (hole > top && comp(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

template <class T, class C>
__host__ __device__ inline void adjust_heap_(T* first, int hole, int len, T value, C comp)
{
    const int top = hole;
    int second = hole;
    while (second < (len - 1) / 2) {
        second = 2 * (second + 1);
        if (comp(first[second], first[second - 1]))
            second--;
        first[hole] = first[second];
        hole = second;
    }
    if ((len & 1) == 0 && second == (len - 2) / 2) {
        second = 2 * (second + 1);
        first[hole] = first[second - 1];
        hole = second - 1;
    }
    push_heap_(first, hole, top, value, comp);
}

template <class T, class C>
__host__ __device__ inline void heap_sort_(T* first, T* last, C comp)
{
    // __partial_sort(first, last, last): __heap_select (make_heap, empty tail loop) + __sort_heap
    int len = (int)(last - first);
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
    while (last - first > 1) {
        --last;
        T value = *last;
        *last = *first;
        adjust_heap_(first, 0, (int)(last - first), value, comp);
    }
}

template <class T, class C>
__host__ __device__ inline void move_median_to_first_(T* result, T* a, T* b, T* c, C comp)
{
    if (comp(*a, *b)) {
        if (comp(*b, *c))
            isort_swap(*result, *b);
        else if (comp(*a, *c))
            isort_swap(*result, *c);
        else
            isort_swap(*result, *a);
    } else if (comp(*a, *c))
        isort_swap(*result, *a);
    else if (comp(*b, *c))
        isort_swap(*result, *c);
    else
        isort_swap(*result, *b);
}

template <class T, class C>
__host__ __device__ inline T* unguarded_partition_(T* first, T* last, T* pivot, C comp)
{
    while (true) {
        while (comp(*first, *pivot))
            ++first;
        --last;
        while (comp(*pivot, *last))
            --last;
        if (!(first < last))
            return first;
        isort_swap(*first, *last);
        ++first;
    }
}

template <class T, class C>
__host__ __device__ inline void unguarded_linear_insert_(T* last, C comp)
{
    T val = *last;
    T* next = last - 1;
    while (comp(val, *next)) {
        *last = *next;
        last = next;
        --next;
    }
    *last = val;
}

template <class T, class C>
__host__ __device__ inline void insertion_sort_(T* first, T* last, C comp)
{
    if (first == last)
        return;
    for (T* i = first + 1; i != last; ++i) {
        if (comp(*i, *first)) {
            T val = *i;
            for (T* p = i; p != first; --p)
                *p = *(p - 1);
            *first = val;
        } else
            unguarded_linear_insert_(i, comp);
    }
}

// std::sort(first, last, comp) — iterative form of __introsort_loop (the recursion on the
// right part becomes an explicit stack; the processing order of sub-ranges is preserved,
// which does not matter for the result since sub-ranges are disjoint).
template <class T, class C>
__host__ __device__ inline void std_sort(T* first, T* last, C comp)
{
    if (first == last)
        return;
    int n = (int)(last - first);
    int lg = 0;
    for (int t = n; t > 1; t >>= 1)
        lg++;
    struct Range {
        T *f, *l;
        int depth;
    };
    Range stack[64];
    int sp = 0;
    stack[sp++] = Range { first, last, lg * 2 };
    while (sp > 0) {
        Range r = stack[--sp];
        T* f = r.f;
        T* l = r.l;
        int depth = r.depth;
        while (l - f > 16) {
            if (depth == 0) {
                heap_sort_(f, l, comp);
                break;
            }
            --depth;
            T* mid = f + (l - f) / 2;
            move_median_to_first_(f, f + 1, mid, l - 1, comp);
            T* cut = unguarded_partition_(f + 1, l, f, comp);
            if (sp < 64)
                stack[sp++] = Range { cut, l, depth };
            l = cut;
        }
    }
    if (last - first > 16) {
        insertion_sort_(first, first + 16, comp);
        for (T* i = first + 16; i != last; ++i)
            unguarded_linear_insert_(i, comp);
    } else
        insertion_sort_(first, last, comp);
}

} // namespace rgbdfe
