// FAST-9/16 arithmetic shared by the per-cell detector of ORBextractor (orb_image.cu) and the whole-level detector
// of the cv::ORB branch (orb_cv.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rgbdfe {

// FAST-9/16 corner strength S: the pixel is a corner at threshold t iff S > t,
// and OpenCV's response (cornerScore<16>) is S - 1.  Returns 0 when S <= th_lo.
// S = max over the 16 arcs of 9 ring pixels of the arc minimum of e[k], with e[k] = a_k - v for a bright corner
// and v - a_k for a dark one.  Both signs ride in the two signed 16-bit lanes of one register (a pixel cannot be
// both: two 9-arcs of a 16-ring overlap), so no sign decision and no ring masks are needed: 16 three-input
// minima give the 3-arcs, 16 more the 9-arcs, 8 three-input maxima the best arc (VIMNMX3.S16x2).
__device__ __forceinline__ int fast_strength(const uint8_t* __restrict__ p, int pitch, int th_lo)
{
    const unsigned v = p[0];
    const unsigned c = ((0u - v) & 0xffffu) | v << 16; // lanes (-v, +v)
    unsigned e[16];
    // a * 0xffff0001 = lanes (a, -a); the lane-wise add of c gives (a - v, v - a)
#define RING(k, off) e[k] = __vadd2((unsigned)p[off] * 0xffff0001u, c)
    RING(0, 3 * pitch);
    RING(1, 3 * pitch + 1);
    RING(2, 2 * pitch + 2);
    RING(3, pitch + 3);
    RING(4, 3);
    RING(5, -pitch + 3);
    RING(6, -2 * pitch + 2);
    RING(7, -3 * pitch + 1);
    RING(8, -3 * pitch);
    RING(9, -3 * pitch - 1);
    RING(10, -2 * pitch - 2);
    RING(11, -pitch - 3);
    RING(12, -3);
    RING(13, pitch - 3);
    RING(14, 2 * pitch - 2);
    RING(15, 3 * pitch - 1);
#undef RING
    unsigned m3[16], m9[16];
#pragma unroll
    for (int k = 0; k < 16; k++)
        m3[k] = __vimin3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++)
        m9[k] = __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
    unsigned best = __vimax3_s16x2(m9[0], m9[1], m9[2]);
#pragma unroll
    for (int k = 3; k < 15; k += 2)
        best = __vimax3_s16x2(best, m9[k], m9[k + 1]);
    best = __vmaxs2(best, m9[15]);
    const int S = max((int)(short)(best & 0xffffu), (int)best >> 16);
    return S > th_lo ? S : 0;
}

// Necessary condition for a 9-arc at threshold t: every opposite pair (k, k+8) of the ring has a member inside
// the arc, so for four of the pairs at least one member must be brighter than v + t (bright corner) or darker
// than v - t (dark corner): min over the pairs of max(a_k, a_k+8) > v + t, or max over the pairs of
// min(a_k, a_k+8) < v - t.  Evaluated for two horizontally adjacent pixels at once in 16-bit lanes
// (VIMNMX.U16x2); the comparisons are additions that carry into bit 15 of each lane.  c1 = 0x7fff7fff - (t, t).
__device__ __forceinline__ unsigned fast_candidate2(unsigned v, unsigned a0, unsigned a8, unsigned a4, unsigned a12, unsigned a2, unsigned a10,
    unsigned a6, unsigned a14, unsigned c1)
{
    const unsigned mb = __vminu2(__vminu2(__vmaxu2(a0, a8), __vmaxu2(a4, a12)), __vminu2(__vmaxu2(a2, a10), __vmaxu2(a6, a14)));
    const unsigned md = __vmaxu2(__vmaxu2(__vminu2(a0, a8), __vminu2(a4, a12)), __vmaxu2(__vminu2(a2, a10), __vminu2(a6, a14)));
    const unsigned tb = mb + (c1 - v); // lane >= 0x8000  <=>  mb > v + t   (no lane ever borrows or carries out)
    const unsigned td = (v + c1) - md; // lane >= 0x8000  <=>  md < v - t
    return (tb | td) & 0x80008000u;
}
// bytes (0,1) / (2,3) of u as two zero-extended 16-bit lanes
__device__ __forceinline__ unsigned lanes_lo(unsigned u) { return __byte_perm(u, 0u, 0x4140); }
__device__ __forceinline__ unsigned lanes_hi(unsigned u) { return __byte_perm(u, 0u, 0x4342); }

} // namespace rgbdfe
