// Image-side ORB kernels (integer, HBM/L2-bound):
//   K1 k_pyramid_level  — ORBextractor::ComputePyramid (Features/ORBextractor.cpp:773-797):
//        level 0 = copy + REFLECT_101 ring; level l = cv::resize(INTER_LINEAR) of
//        level l-1 (8-bit fixed point, 11-bit coefficients) + ring, in one pass.
//   K5 k_gauss7         — cv::GaussianBlur(7x7, sigma 2, REFLECT_101) call site :745-746,
//        fixed-point kernel {18,34,48,56,48,34,18}/256, one rounding (>>16).
//   K2 k_fast_cells     — per-30px-cell cv::FAST(th=ini, nonmax) with fallback to th=min
//        (:636-672): FAST-9/16 corner strength, cell-local 3x3 NMS, row-major output.
// blockIdx.z (or .y) is the frame slot: every kernel processes a batch of frames.
#include "common.cuh"
#include "fast_core.cuh"
#include "tma.cuh"

#include <memory>
#include <string>
#include <vector>

namespace rgbdfe {

// ------------------------------------------------------------------ K1 ----
// Level 0: copy + REFLECT_101 ring.  One thread per 4 horizontally adjacent padded pixels (one 32-bit store);
// interior words are two aligned loads and a funnel shift, ring words go byte by byte.  channels == 3: the source
// is the BGR colour image and Frame's cvtColor(BGR2GRAY) (Core/Frame.cpp:47) is applied on the way in -- OpenCV's
// 8-bit fixed point: (B * 3735 + G * 19235 + R * 9798 + 2^14) >> 15.
__device__ __forceinline__ unsigned bgr2gray(unsigned b, unsigned g, unsigned r) { return (b * 3735u + g * 19235u + r * 9798u + 16384u) >> 15; }

__global__ void __launch_bounds__(256) k_pyramid_base(int w, int h, int pstride, unsigned pyr_off, const __grid_constant__ ImgPtrs img, int img_stride,
    int channels, uint8_t* __restrict__ pyr, size_t pyr_slot)
{
    const int pw = w + 2 * EDGE_THRESHOLD, ph = h + 2 * EDGE_THRESHOLD;
    const int px0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px0 >= pw || py >= ph)
        return;
    const int slot = blockIdx.z;
    const uint8_t* src = static_cast<const uint8_t*>(img.p[slot]) + (size_t)reflect101(py - EDGE_THRESHOLD, h) * img_stride;
    const bool interior = px0 >= EDGE_THRESHOLD && px0 + 3 < EDGE_THRESHOLD + w;
    unsigned packed = 0;
    if (channels == 3) {
        const uint8_t* q = src + 3 * (px0 - EDGE_THRESHOLD);
        if (interior && (reinterpret_cast<size_t>(q) & 3) == 0) { // 4 pixels = 3 aligned words
            const unsigned w0 = reinterpret_cast<const unsigned*>(q)[0], w1 = reinterpret_cast<const unsigned*>(q)[1],
                           w2 = reinterpret_cast<const unsigned*>(q)[2];
            packed = bgr2gray(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u) | bgr2gray(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u) << 8
                | bgr2gray((w1 >> 16) & 255u, w1 >> 24, w2 & 255u) << 16 | bgr2gray((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24) << 24;
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint8_t* p = src + 3 * reflect101(min(px0 + k, pw - 1) - EDGE_THRESHOLD, w);
                packed |= bgr2gray(p[0], p[1], p[2]) << (8 * k);
            }
        }
    } else if (interior) {
        const uint8_t* q = src + (px0 - EDGE_THRESHOLD);
        const unsigned mis = (unsigned)(reinterpret_cast<size_t>(q) & 3);
        const unsigned* qa = reinterpret_cast<const unsigned*>(q - mis);
        packed = __funnelshift_r(qa[0], mis ? qa[1] : 0u, 8 * mis); // qa[1] is touched only when it holds wanted bytes
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            packed |= (unsigned)src[reflect101(min(px0 + k, pw - 1) - EDGE_THRESHOLD, w)] << (8 * k);
    }
    *reinterpret_cast<unsigned*>(pyr + slot * pyr_slot + pyr_off + (size_t)py * pstride + px0) = packed; // pstride is a multiple of 64
}

// Level 0 of a gray image, 16 padded pixels per thread (one 16-byte store): threads whose span is interior read five
// aligned words and funnel-shift them into place (the source row starts at any byte offset); spans that touch the
// ring go byte by byte through REFLECT_101.
__global__ void __launch_bounds__(256) k_pyramid_base16(int w, int h, int pstride, unsigned pyr_off, const __grid_constant__ ImgPtrs img, int img_stride,
    uint8_t* __restrict__ pyr, size_t pyr_slot)
{
    const int px0 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px0 >= pstride || py >= h + 2 * EDGE_THRESHOLD)
        return;
    const int pw = w + 2 * EDGE_THRESHOLD;
    const int slot = blockIdx.z;
    const uint8_t* src = static_cast<const uint8_t*>(img.p[slot]) + (size_t)reflect101(py - EDGE_THRESHOLD, h) * img_stride;
    uint4 out;
    if (px0 >= EDGE_THRESHOLD && px0 + 15 < EDGE_THRESHOLD + w) {
        const uint8_t* q = src + (px0 - EDGE_THRESHOLD);
        const unsigned mis = (unsigned)(reinterpret_cast<size_t>(q) & 3);
        const unsigned* qa = reinterpret_cast<const unsigned*>(q - mis);
        const unsigned w0 = qa[0], w1 = qa[1], w2 = qa[2], w3 = qa[3], w4 = mis ? qa[4] : 0u; // qa[4] is touched only when it holds wanted bytes
        const unsigned sh = 8 * mis;
        out = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
    } else {
        unsigned o[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            unsigned packed = 0;
#pragma unroll
            for (int k = 0; k < 4; k++)
                packed |= (unsigned)src[reflect101(min(px0 + 4 * i + k, pw - 1) - EDGE_THRESHOLD, w)] << (8 * k);
            o[i] = packed;
        }
        out = make_uint4(o[0], o[1], o[2], o[3]);
    }
    *reinterpret_cast<uint4*>(pyr + slot * pyr_slot + pyr_off + (size_t)py * pstride + px0) = out; // pstride is a multiple of 64
}

// Level l >= 1: cv::resize(INTER_LINEAR) of level l-1 (8-bit fixed point, 11-bit coefficients, one rounding per
// pass as in OpenCV's resize.cpp) + ring in one pass over the PADDED level: ring pixels are resampled at their
// reflected coordinates, which is what copyMakeBorder would copy.  The tables are indexed by padded coordinates
// (ring reflection, edge clamps and the pitch padding resolved on the host), so the kernel has no border arithmetic.
// One thread = 4 horizontally adjacent padded pixels x PYR_ROWS rows.  The host lists, per group of PYR_ROWS rows,
// the distinct source rows the group reads (6 at scale 1.2 instead of 2 per output row) and which two of them each
// output row blends: the thread blends every listed source row horizontally first -- all its byte loads are
// independent and in flight together -- parks the 16-bit results in its own shared-memory column (registers cannot
// be indexed by the row number) and then blends vertically.  No barrier: a thread reads back only what it wrote.
// The source is read where it lies (level l-1 was written by the previous launch and sits in L2; neighbouring
// threads share its sectors through L1).
struct PyrLevel {
    int ph, pstride;             // this level: padded rows, row pitch
    unsigned pyr_off;
    unsigned pyr_off_src;        // level l-1
    int tab_x_off, tab_g_off;
};
template <int NSRC>
__global__ void __launch_bounds__(256) k_pyramid_resize_direct(const __grid_constant__ PyrLevel V, uint8_t* __restrict__ pyr, size_t pyr_slot,
    const int2* __restrict__ xpad, const PyrGroup* __restrict__ groups)
{
    __shared__ uint2 s_t[NSRC][256];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int px0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int g = blockIdx.y * blockDim.y + threadIdx.y;
    const int py0 = g * PYR_ROWS;
    if (px0 >= V.pstride || py0 >= V.ph)
        return;
    uint8_t* slot_base = pyr + blockIdx.z * pyr_slot;
    const uint8_t* src = slot_base + V.pyr_off_src;
    const int4* xp = reinterpret_cast<const int4*>(xpad + V.tab_x_off + px0);
    const int4 e0 = xp[0], e1 = xp[1];
    const int4* gp = reinterpret_cast<const int4*>(groups + V.tab_g_off + g);
    const int4 o0 = gp[0], o1 = gp[1], bt = gp[2], m = gp[3];
    const int off[8] = { o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w };
    const int sx[4] = { e0.x, e0.z, e1.x, e1.z };
    const unsigned al[4] = { (unsigned)e0.y, (unsigned)e0.w, (unsigned)e1.y, (unsigned)e1.w };
#pragma unroll
    for (int s = 0; s < NSRC; s++) {
        const uint8_t* row = src + off[s];
        unsigned t[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint8_t* q = row + sx[k];
            t[k] = (q[0] * (al[k] & 0xffffu) + q[1] * (al[k] >> 16)) >> 4; // <= 255 * 2049 >> 4: fits 16 bits
        }
        s_t[s][tid] = make_uint2(t[1] * 65536u + t[0], t[3] * 65536u + t[2]);
    }
    const unsigned beta[4] = { (unsigned)bt.x, (unsigned)bt.y, (unsigned)bt.z, (unsigned)bt.w };
    uint8_t* dst = slot_base + V.pyr_off + (size_t)py0 * V.pstride + px0;
#pragma unroll
    for (int j = 0; j < PYR_ROWS; j++) {
        if (py0 + j < V.ph) {
            const unsigned sel = (unsigned)m.x >> (8 * j);
            const uint2 a = s_t[sel & 15u][tid], b = s_t[(sel >> 4) & 15u][tid];
            const int b0 = (int)(beta[j] & 0xffffu), b1 = (int)(beta[j] >> 16);
            // cv's saturate_cast never binds here: both coefficient pairs sum to <= 2049, so the value is <= 255
            const unsigned v0 = (unsigned)((((b0 * (int)(a.x & 0xffffu)) >> 16) + ((b1 * (int)(b.x & 0xffffu)) >> 16) + 2) >> 2);
            const unsigned v1 = (unsigned)((((b0 * (int)(a.x >> 16)) >> 16) + ((b1 * (int)(b.x >> 16)) >> 16) + 2) >> 2);
            const unsigned v2 = (unsigned)((((b0 * (int)(a.y & 0xffffu)) >> 16) + ((b1 * (int)(b.y & 0xffffu)) >> 16) + 2) >> 2);
            const unsigned v3 = (unsigned)((((b0 * (int)(a.y >> 16)) >> 16) + ((b1 * (int)(b.y >> 16)) >> 16) + 2) >> 2);
            *reinterpret_cast<unsigned*>(dst + (size_t)j * V.pstride) = ((v3 * 256u + v2) * 256u + v1) * 256u + v0; // bytes past the padded width are pitch padding
        }
    }
}

// ------------------------------------------------------------------ K5 ----
// Separable 7-tap blur with the 8-bit fixed-point kernel {18,34,48,56,48,34,18}/256 on packed bytes:
// the horizontal pass is two IDP4A per output (4 window bytes x 4 coefficient bytes each), the vertical
// pass four IDP2A per output (two vertically adjacent 16-bit row sums x two coefficient bytes).  The
// level's 19 px REFLECT_101 ring in the pyramid is exactly the border GaussianBlur asks for, so the
// tile is staged straight from the padded image with 16-byte loads and no border arithmetic.
#define BLUR_TW 64
#define BLUR_TH 32
#define BLUR_SROWS (BLUR_TH + 6)
#define BLUR_SPITCH 80 // staged bytes per row: level columns tx0-3 .. tx0+76
// Geometry of the blur's tile grid, passed by value (constant bank): the prologue has no dependent global loads.
struct BlurLevel {
    int w, h, pstride, bstride;
    unsigned pyr_off, blur_off;
};
struct BlurGrid {
    int nlevels;
    int tile_start[RGBDFE_MAX_LEVELS + 1];
    BlurLevel lv[RGBDFE_MAX_LEVELS];
};
__global__ void __launch_bounds__(256) k_gauss7(const __grid_constant__ BlurGrid G, const uint8_t* __restrict__ pyr, size_t pyr_slot,
    uint8_t* __restrict__ blur, size_t blur_slot)
{
    // blockIdx.x enumerates tiles of all levels
    int level = 0;
    while (level + 1 < G.nlevels && (int)blockIdx.x >= G.tile_start[level + 1])
        level++;
    const BlurLevel L = G.lv[level];
    const int t = blockIdx.x - G.tile_start[level];
    const int tiles_x = (L.w + BLUR_TW - 1) / BLUR_TW;
    const int ty = t / tiles_x;
    const int tx0 = (t - ty * tiles_x) * BLUR_TW, ty0 = ty * BLUR_TH;
    const int slot = blockIdx.y;
    // level pixel (x, y) sits at padded (x + 19, y + 19); `base` is level pixel (-3, -3): 16-byte aligned
    // because pyr_off and pstride are multiples of 64 and EDGE_THRESHOLD - 3 == 16
    const uint8_t* base = pyr + slot * pyr_slot + L.pyr_off + (size_t)(EDGE_THRESHOLD - 3) * L.pstride + (EDGE_THRESHOLD - 3);

    __shared__ __align__(16) uint8_t s_src[BLUR_SROWS][BLUR_SPITCH];
    __shared__ __align__(16) unsigned short s_h[BLUR_SROWS + 2][BLUR_TW]; // rows 38, 39 only ever meet a zero coefficient
    const int tid = threadIdx.x;
    for (int i = tid; i < BLUR_SROWS * (BLUR_SPITCH / 16); i += 256) {
        const int r = i / (BLUR_SPITCH / 16), q = i - r * (BLUR_SPITCH / 16);
        const int y = min(ty0 + r, L.h + 5); // level row ty0 + r - 3, clamped to the last row a valid output reads
        *reinterpret_cast<uint4*>(&s_src[r][16 * q]) = *reinterpret_cast<const uint4*>(base + (size_t)y * L.pstride + tx0 + 16 * q);
    }
    if (tid < BLUR_TW)
        *reinterpret_cast<unsigned*>(&s_h[BLUR_SROWS][2 * tid]) = 0u;
    __syncthreads();
    // horizontal: 4 adjacent outputs per thread from 12 staged bytes
    constexpr unsigned KH0 = 18u | 34u << 8 | 48u << 16 | 56u << 24, KH1 = 48u | 34u << 8 | 18u << 16;
    for (int i = tid; i < BLUR_SROWS * (BLUR_TW / 4); i += 256) {
        const int r = i >> 4, k = i & 15;
        const unsigned* row = reinterpret_cast<const unsigned*>(&s_src[r][0]);
        const unsigned w0 = row[k], w1 = row[k + 1], w2 = row[k + 2];
        const unsigned a0 = __dp4a(w0, KH0, __dp4a(w1, KH1, 0u));
        const unsigned a1 = __dp4a(__funnelshift_r(w0, w1, 8), KH0, __dp4a(__funnelshift_r(w1, w2, 8), KH1, 0u));
        const unsigned a2 = __dp4a(__funnelshift_r(w0, w1, 16), KH0, __dp4a(__funnelshift_r(w1, w2, 16), KH1, 0u));
        const unsigned a3 = __dp4a(__funnelshift_r(w0, w1, 24), KH0, __dp4a(__funnelshift_r(w1, w2, 24), KH1, 0u));
        *reinterpret_cast<uint2*>(&s_h[r][4 * k]) = make_uint2(a0 | a1 << 16, a2 | a3 << 16); // sums are <= 255 * 256
    }
    __syncthreads();
    // vertical: one thread = 2 adjacent columns x the 4 output rows of one parity in an 8-row band, so that the
    // (row, row + 1) pairs it packs for IDP2A serve all four of its outputs
    constexpr unsigned KV0 = KH0, KV1 = KH1; // lo half (18,34) / hi half (48,56); lo half (48,34) / hi half (18,0)
    const int cp = tid & 31, par = (tid >> 5) & 1, band = tid >> 6;
    const int r0 = band * 8 + par;
    unsigned lo[7], hi[7];
#pragma unroll
    for (int j = 0; j < 7; j++) {
        const unsigned ra = *reinterpret_cast<const unsigned*>(&s_h[r0 + 2 * j][2 * cp]);
        const unsigned rb = *reinterpret_cast<const unsigned*>(&s_h[r0 + 2 * j + 1][2 * cp]);
        lo[j] = __byte_perm(ra, rb, 0x5410); // column 2cp:     (row, row + 1)
        hi[j] = __byte_perm(ra, rb, 0x7632); // column 2cp + 1
    }
    uint8_t* dst = blur + slot * blur_slot + L.blur_off;
    const int x = tx0 + 2 * cp;
#pragma unroll
    for (int m = 0; m < 4; m++) {
        const int y = ty0 + r0 + 2 * m;
        // one rounding at the end: (sum + 2^15) >> 16; the sum cannot exceed 255.5 * 2^16, no saturation needed
        const unsigned va = __dp2a_hi(lo[m + 3], KV1, __dp2a_lo(lo[m + 2], KV1, __dp2a_hi(lo[m + 1], KV0, __dp2a_lo(lo[m], KV0, 32768u))));
        const unsigned vb = __dp2a_hi(hi[m + 3], KV1, __dp2a_lo(hi[m + 2], KV1, __dp2a_hi(hi[m + 1], KV0, __dp2a_lo(hi[m], KV0, 32768u))));
        if (y < L.h && x < L.w) // x + 1 == L.w only for odd widths, where the row pitch has padding
            *reinterpret_cast<unsigned short*>(dst + (size_t)y * L.bstride + x) = (unsigned short)((va >> 16) | ((vb >> 16) << 8));
    }
}

// ------------------------------------------------------------------ K2 ----
// Geometry of the FAST cell grid of every level, passed by value (constant bank): the kernel's prologue has no
// dependent global loads.
struct FastGrid {
    int iniTh, minTh;
    int rows;     // rows of the largest cell window (max hCell + 6): the shared-memory images are sized for the configuration, not for CELL_MAX_DIM
    int cand_cap; // pixels of the largest cell window
    unsigned inv_small[32]; // 0xffffffff / d + 1 for d = 1..31: i / d == umulhi(i, inv_small[d]) (groups per cell row)
};
// Tensor maps of the pyramid levels, resident in global memory (one 128-byte descriptor per level): the padded level of
// every frame slot as a (pstride, h + 38, slots) u8 tensor, box = (96, hCell + 6, 1).
struct FastMaps {
    CUtensorMap map[RGBDFE_MAX_LEVELS];
};
// One FAST cell, precomputed on the host (ORBextractor.cpp:636-653): window origin and size in level coordinates
// (minBorder-relative), origin of the cell in the output coordinates, level; cw == 0 marks a skipped cell.
struct FastCell {
    short iniX, iniY, cw, ch;
    short outX, outY, level, box_rows; // box_rows = hCell + 6: rows of the level's TMA box
};

#define FAST_THREADS 128 // 4 warps per cell: 225 pixel groups of the pair test in two passes, the later (sparser) phases idle fewer warps; measured 0.225 / 0.196 / 0.207 ms at 256 / 128 / 64
__global__ void __launch_bounds__(FAST_THREADS) k_fast_cells(const __grid_constant__ FastGrid G, const FastMaps* __restrict__ maps, const FastCell* __restrict__ cells,
    int* __restrict__ cell_count, unsigned* __restrict__ cell_list, int cell_slot)
{
    const FastCell cd = cells[blockIdx.x];
    const int slot = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31;
    int* out_count = cell_count + (size_t)slot * cell_slot + blockIdx.x;
    unsigned* out_list = cell_list + ((size_t)slot * cell_slot + blockIdx.x) * CELL_CAP;
    const int cw = cd.cw, ch = cd.ch;
    if (cw == 0) {
        if (tid == 0)
            *out_count = 0;
        return;
    }
    constexpr int PITCH = CELL_MAX_DIM + 24; // window + up to 15 bytes of alignment shift, multiple of 16
    // dynamic shared memory sized for this configuration's largest cell (a 37 x 37 window at 640 x 480: 12.6 KB instead of the
    // 34.5 KB of a CELL_MAX_DIM window, so that threads, not shared memory, limit the CTAs per SM)
    extern __shared__ __align__(128) uint8_t fast_smem[];
    uint8_t* const s_img = fast_smem;                                  // [rows][PITCH], TMA destination
    uint8_t* const s_S = fast_smem + (size_t)G.rows * PITCH;           // [rows][PITCH] corner strengths (PITCH * rows is a multiple of 32)
    unsigned short* const s_cand = reinterpret_cast<unsigned short*>(s_S + (size_t)G.rows * PITCH); // pixels that pass the pair test: r << 8 | c
    unsigned short* const s_keep = s_cand + G.cand_cap;               // NMS survivors of this attempt: r << 8 | c, unordered
#define S_IMG(r, x) s_img[(r) * PITCH + (x)]
#define S_S(r, x) s_S[(r) * PITCH + (x)]
    __shared__ int s_ncand, s_n;
    __shared__ __align__(8) uint64_t s_bar;
    // The cell's window is staged by the TMA unit: one thread arms the barrier and issues the tensor copy (box 96 x
    // (hCell + 6) bytes; the box must start at a 16-byte aligned address, so it starts up to 15 columns left of the
    // window: image column c lives at s_img[r][c + mis]; the 19 px ring keeps every window inside the level), the other
    // threads clear the strength map meanwhile.
    const int wx = EDGE_THRESHOLD + cd.iniX, mis = wx & 15;
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_expect_tx(&s_bar, (unsigned)PITCH * (unsigned)cd.box_rows);
        tma_load_3d(s_img, &maps->map[cd.level], wx - mis, EDGE_THRESHOLD + cd.iniY, slot, &s_bar);
        s_ncand = 0;
        s_n = 0;
    }
    for (int i = tid; i < (PITCH / 16) * ch; i += FAST_THREADS) // 16-byte stores: six per row, one or two passes
        reinterpret_cast<uint4*>(s_S)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    mbar_wait(&s_bar, 0);
    // image column c lives at s_img[r][c + mis]; s_S uses the same shift
    const int dw = cw - 6, dh = ch - 6;
    const unsigned lt = (1u << lane) - 1;
    // pair test geometry: 4 adjacent pixels per thread = one word-aligned group of shared-memory columns [4g, 4g + 4)
    const int s_first = 3 + mis; // shared-memory column of the first tested pixel of a row
    const int g_lo = s_first >> 2, gpr = ((cw - 4 + mis) >> 2) - g_lo + 1;
    const unsigned inv_gpr = G.inv_small[gpr];
    const int ngroups = gpr * dh;
    // cv::FAST(iniThFAST) first; only a cell that comes out EMPTY is redone at minThFAST (:655-661).  Everything -- the pair
    // test, the strengths, the NMS -- runs at the attempt's own threshold: a neighbour counts in the NMS only if it is a
    // corner at that threshold, i.e. a candidate of the same attempt, and the candidates at 20 are a small part of those at 7.
    int th = G.iniTh;
    int ncand = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        const unsigned c1 = 0x7fff7fffu - ((unsigned)th | (unsigned)th << 16);
        for (int i0 = (tid & ~31); i0 < ngroups; i0 += FAST_THREADS) { // warp-uniform trip count
            const int i = i0 + lane;
            unsigned rA = 0, rB = 0;
            int r = 0, s0 = 0;
            if (i < ngroups) {
                const int rr = (int)__umulhi((unsigned)i, inv_gpr);
                const int g = g_lo + (i - rr * gpr);
                r = 3 + rr;
                s0 = 4 * g;
                const unsigned* w = reinterpret_cast<const unsigned*>(&S_IMG(r, 0)) + g; // PITCH is a multiple of 4
                constexpr int RW = PITCH / 4;
                const unsigned cm = w[-1], cc = w[0], cp = w[1];
                const unsigned up3 = w[-3 * RW], dn3 = w[3 * RW];
                const unsigned um = w[-2 * RW - 1], uc = w[-2 * RW], upn = w[-2 * RW + 1];
                const unsigned dm = w[2 * RW - 1], dc = w[2 * RW], dpn = w[2 * RW + 1];
                const unsigned u4 = __funnelshift_r(cc, cp, 24), u12 = __funnelshift_r(cm, cc, 8); // columns +3 / -3 of the centre row
                const unsigned u2 = __funnelshift_r(dc, dpn, 16), u14 = __funnelshift_r(dm, dc, 16); // row +2: columns +2 / -2
                const unsigned u6 = __funnelshift_r(uc, upn, 16), u10 = __funnelshift_r(um, uc, 16); // row -2: columns +2 / -2
                rA = fast_candidate2(lanes_lo(cc), lanes_lo(dn3), lanes_lo(up3), lanes_lo(u4), lanes_lo(u12), lanes_lo(u2), lanes_lo(u10),
                    lanes_lo(u6), lanes_lo(u14), c1);
                rB = fast_candidate2(lanes_hi(cc), lanes_hi(dn3), lanes_hi(up3), lanes_hi(u4), lanes_hi(u12), lanes_hi(u2), lanes_hi(u10),
                    lanes_hi(u6), lanes_hi(u14), c1);
            }
            const int rel = s0 - s_first; // tested pixels of the row are rel + j in [0, dw)
            const bool k0 = (rA & 0x8000u) && (unsigned)rel < (unsigned)dw, k1 = (rA >> 31) && (unsigned)(rel + 1) < (unsigned)dw;
            const bool k2 = (rB & 0x8000u) && (unsigned)(rel + 2) < (unsigned)dw, k3 = (rB >> 31) && (unsigned)(rel + 3) < (unsigned)dw;
            if (__any_sync(0xffffffffu, k0 | k1 | k2 | k3)) { // most groups of 128 pixels hold no candidate at all
                const unsigned b0 = __ballot_sync(0xffffffffu, k0), b1 = __ballot_sync(0xffffffffu, k1);
                const unsigned b2 = __ballot_sync(0xffffffffu, k2), b3 = __ballot_sync(0xffffffffu, k3);
                const int n0 = __popc(b0), n1 = __popc(b1), n2 = __popc(b2), n3 = __popc(b3);
                int base = 0;
                if (lane == 0)
                    base = atomicAdd(&s_ncand, n0 + n1 + n2 + n3);
                base = __shfl_sync(0xffffffffu, base, 0);
                const int key = r << 8 | (s0 - mis); // window column of pixel 0 of the group
                if (k0)
                    s_cand[base + __popc(b0 & lt)] = (unsigned short)key;
                if (k1)
                    s_cand[base + n0 + __popc(b1 & lt)] = (unsigned short)(key + 1);
                if (k2)
                    s_cand[base + n0 + n1 + __popc(b2 & lt)] = (unsigned short)(key + 2);
                if (k3)
                    s_cand[base + n0 + n1 + n2 + __popc(b3 & lt)] = (unsigned short)(key + 3);
            }
        }
        __syncthreads();
        ncand = s_ncand;
        for (int i = tid; i < ncand; i += FAST_THREADS) {
            const int r = s_cand[i] >> 8, c = s_cand[i] & 255;
            const int S = fast_strength(&S_IMG(r, c + mis), PITCH, th);
            S_S(r, c + mis) = (uint8_t)min(S, 255);
        }
        __syncthreads();
        // 3x3 non-max suppression inside the cell's window (cv::FAST, nonmaxSuppression = true), candidates only
        for (int i0 = 0; i0 < ncand; i0 += FAST_THREADS) {
            const int i = i0 + tid;
            bool keep = false;
            int rc = 0;
            if (i < ncand) {
                rc = s_cand[i];
                const int r = rc >> 8, c = (rc & 255) + mis;
                const int S = S_S(r, c);
                if (S > th) {
                    const int sc = S - 1;
                    keep = true;
#pragma unroll
                    for (int dy = -1; dy <= 1; dy++)
#pragma unroll
                        for (int dx = -1; dx <= 1; dx++) {
                            if (dx == 0 && dy == 0)
                                continue;
                            const int Sn = S_S(r + dy, c + dx);
                            const int sn = Sn > th ? Sn - 1 : 0;
                            keep = keep && (sc > sn);
                        }
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, keep);
            if (bal) {
                int base = 0;
                if (lane == 0)
                    base = atomicAdd(&s_n, __popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (keep)
                    s_keep[base + __popc(bal & lt)] = (unsigned short)rc;
            }
        }
        __syncthreads();
        if (s_n > 0 || th == G.minTh)
            break;
        th = G.minTh; // cv::FAST returned nothing at iniThFAST: the cell is redone at minThFAST (:658-661)
        if (tid == 0)
            s_ncand = 0; // the candidates at the lower threshold include every one above; their strengths are rewritten
        __syncthreads();
    }
    // row-major output order (cv::FAST emits keypoints row by row): rank each survivor by its (r, c) key
    const int n = s_n;
    for (int i = tid; i < n; i += FAST_THREADS) {
        const int key = s_keep[i];
        int rank = 0;
        for (int j = 0; j < n; j++)
            rank += s_keep[j] < key;
        if (rank < CELL_CAP) {
            const int r = key >> 8, c = key & 255;
            out_list[rank] = pack_cand(c + cd.outX, r + cd.outY, S_S(r, c + mis) - 1);
        }
    }
    if (tid == 0)
        *out_count = min(n, CELL_CAP);
}

// ---------------------------------------------------------------- launchers
int launch_pyramid(rgbdfe_ctx* c, int count, const ImgPtrs& gray, int gray_stride, int channels)
{
    const OrbConfig& cfg = c->cfg;
    ProfScope prof(c, KID_PYRAMID);
    for (int l = 0; l < cfg.nlevels; l++) {
        const LevelDesc& L = cfg.lv[l];
        const int pw = L.w + 2 * EDGE_THRESHOLD, ph = L.h + 2 * EDGE_THRESHOLD;
        if (l == 0) {
            if (channels == 1) {
                dim3 blk(16, 16), grd((L.pstride / 16 + 15) / 16, (ph + 15) / 16, count);
                k_pyramid_base16<<<grd, blk, 0, c->stream>>>(L.w, L.h, L.pstride, L.pyr_off, gray, gray_stride, c->d_pyr, c->pyr_slot);
            } else {
                dim3 blk(32, 8), grd((pw + 127) / 128, (ph + 7) / 8, count);
                k_pyramid_base<<<grd, blk, 0, c->stream>>>(L.w, L.h, L.pstride, L.pyr_off, gray, gray_stride, channels, c->d_pyr, c->pyr_slot);
            }
        } else {
            const LevelDesc& P = cfg.lv[l - 1];
            PyrLevel V;
            V.ph = ph, V.pstride = L.pstride, V.pyr_off = L.pyr_off, V.pyr_off_src = P.pyr_off;
            V.tab_x_off = L.tab_x_off, V.tab_g_off = L.tab_y_off;
            dim3 blk(64, 4), grd((L.pstride / 4 + 63) / 64, (ph + 4 * PYR_ROWS - 1) / (4 * PYR_ROWS), count);
            if (c->pyr_nsrc <= 6)
                k_pyramid_resize_direct<6><<<grd, blk, 0, c->stream>>>(V, c->d_pyr, c->pyr_slot, c->d_tab_xpad, c->d_tab_group);
            else
                k_pyramid_resize_direct<8><<<grd, blk, 0, c->stream>>>(V, c->d_pyr, c->pyr_slot, c->d_tab_xpad, c->d_tab_group);
        }
        c->launches++;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

int launch_blur(rgbdfe_ctx* c, int count)
{
    const OrbConfig& cfg = c->cfg;
    const int total = cfg.blur_tile_start[cfg.nlevels];
    ProfScope prof(c, KID_BLUR);
    BlurGrid G;
    memset(&G, 0, sizeof(G));
    G.nlevels = cfg.nlevels;
    for (int l = 0; l <= cfg.nlevels; l++)
        G.tile_start[l] = cfg.blur_tile_start[l];
    for (int l = 0; l < cfg.nlevels; l++) {
        const LevelDesc& L = cfg.lv[l];
        G.lv[l] = BlurLevel { L.w, L.h, L.pstride, L.bstride, L.pyr_off, L.blur_off };
    }
    k_gauss7<<<dim3(total, count), 256, 0, c->stream>>>(G, c->d_pyr, c->pyr_slot, c->d_blur, c->blur_slot);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

// Per-configuration state of K2: the cell table and the tensor maps of the pyramid levels (rebuilt when the level
// geometry or the pyramid buffer changes).
static int prepare_fast(rgbdfe_ctx* c)
{
    const OrbConfig& cfg = c->cfg;
    if (c->fast_ready && c->fast_key_pyr == c->d_pyr && c->fast_key_w == cfg.w && c->fast_key_h == cfg.h && c->fast_key_levels == cfg.nlevels
        && c->fast_key_slots == c->max_batch)
        return RGBDFE_OK;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) {
        c->set_error("cuTensorMapEncodeTiled is not available from this driver (TMA staging needs CUDA >= 12)");
        return RGBDFE_E_CUDA;
    }
    FastGrid& G = *static_cast<FastGrid*>(c->fast_grid);
    memset(&G, 0, sizeof(G));
    for (int d = 1; d < 32; d++)
        G.inv_small[d] = 0xffffffffu / (unsigned)d + 1u;
    FastMaps* hm = new FastMaps();
    std::unique_ptr<FastMaps> hm_owner(hm);
    std::vector<FastCell> cells(cfg.total_cells);
    for (int l = 0; l < cfg.nlevels; l++) {
        const LevelDesc& L = cfg.lv[l];
        G.rows = std::max(G.rows, (L.hCell + 6 + 3) & ~3); // multiple of 4: PITCH * rows stays a multiple of 128
        G.cand_cap = std::max(G.cand_cap, ((L.wCell + 6) * (L.hCell + 6) + 7) & ~7);
        for (int ci = 0; ci < L.nRows; ci++)
            for (int cj = 0; cj < L.nCols; cj++) {
                // ORBextractor.cpp:636-653 (coordinates relative to minBorder = EDGE_THRESHOLD - 3 = 16)
                FastCell& f = cells[L.cell_off + ci * L.nCols + cj];
                const int iniY = 16 + ci * L.hCell, iniX = 16 + cj * L.wCell;
                int maxY = iniY + L.hCell + 6, maxX = iniX + L.wCell + 6;
                const bool skip = (iniY >= L.maxBorderY - 3) || (iniX >= L.maxBorderX - 6);
                maxY = std::min(maxY, L.maxBorderY);
                maxX = std::min(maxX, L.maxBorderX);
                const int cw = maxX - iniX, ch = maxY - iniY;
                f.iniX = (short)iniX, f.iniY = (short)iniY;
                f.cw = (short)((skip || cw < 7 || ch < 7) ? 0 : cw), f.ch = (short)ch;
                f.outX = (short)(cj * L.wCell), f.outY = (short)(ci * L.hCell);
                f.level = (short)l, f.box_rows = (short)(L.hCell + 6);
            }
        const cuuint64_t dims[3] = { (cuuint64_t)L.pstride, (cuuint64_t)(L.h + 2 * EDGE_THRESHOLD), (cuuint64_t)c->max_batch };
        const cuuint64_t strides[2] = { (cuuint64_t)L.pstride, (cuuint64_t)c->pyr_slot };
        const cuuint32_t box[3] = { CELL_MAX_DIM + 24, (cuuint32_t)(L.hCell + 6), 1 };
        const cuuint32_t estr[3] = { 1, 1, 1 };
        const CUresult r = enc(&hm->map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, c->d_pyr + L.pyr_off, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            c->set_error("cuTensorMapEncodeTiled failed for a pyramid level (code " + std::to_string((int)r) + ")");
            return RGBDFE_E_CUDA;
        }
    }
    cudaFree(c->d_fast_cells);
    cudaFree(c->d_fast_maps);
    c->d_fast_cells = nullptr, c->d_fast_maps = nullptr;
    CUDA_TRY(c, cudaMalloc(&c->d_fast_maps, sizeof(FastMaps)));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_fast_maps, hm, sizeof(FastMaps), cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMalloc(&c->d_fast_cells, cells.size() * sizeof(FastCell)));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_fast_cells, cells.data(), cells.size() * sizeof(FastCell), cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream)); // `cells` is a local
    c->fast_ready = true;
    c->fast_key_pyr = c->d_pyr, c->fast_key_w = cfg.w, c->fast_key_h = cfg.h, c->fast_key_levels = cfg.nlevels, c->fast_key_slots = c->max_batch;
    return RGBDFE_OK;
}

int launch_fast(rgbdfe_ctx* c, int count)
{
    if (!c->fast_grid)
        c->fast_grid = new FastGrid();
    const int rc = prepare_fast(c);
    if (rc != RGBDFE_OK)
        return rc;
    ProfScope prof(c, KID_FAST);
    FastGrid& G = *static_cast<FastGrid*>(c->fast_grid);
    G.iniTh = c->cfg.iniTh, G.minTh = c->cfg.minTh;
    const size_t smem = 2 * (size_t)G.rows * (CELL_MAX_DIM + 24) + 4 * (size_t)G.cand_cap;
    CUDA_TRY(c, ensure_dyn_smem(c, k_fast_cells, smem));
    k_fast_cells<<<dim3(c->cfg.total_cells, count), FAST_THREADS, smem, c->stream>>>(G, static_cast<const FastMaps*>(c->d_fast_maps), static_cast<const FastCell*>(c->d_fast_cells), c->d_cell_count,
        c->d_cell_list, c->cell_slot);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

void free_fast_state(rgbdfe_ctx* c)
{
    cudaFree(c->d_fast_cells);
    cudaFree(c->d_fast_maps);
    c->d_fast_cells = nullptr, c->d_fast_maps = nullptr;
    delete static_cast<FastGrid*>(c->fast_grid);
    c->fast_grid = nullptr;
    c->fast_ready = false;
}

} // namespace rgbdfe
