// Row N3: the `Extractor(ORB, ORB)` branch of the reference = OpenCV's cv::ORB with Harris scoring.
//   mpDetector   = cv::ORB::create(nfeatures, scaleFactor, nlevels)   Features/Extractor.cpp:163-166
//   mpDescriptor = cv::ORB::create()                                  Features/Extractor.cpp:216-218
//   detect -> retainBest(nfeatures) -> compute                        Features/Extractor.cpp:55-59
// The arithmetic is OpenCV's (features2d/src/orb.cpp, imgproc resize.cpp / filter.simd.hpp, keypoint.cpp); the CPU
// restatement pinned against cv2 4.13 is oracle/cvorb.cpp.  Kernels (sm_100a, no FMA except where OpenCV's build fuses):
//   k_cv_pyr_base / k_cv_pyr_level  pyramid chain resize(INTER_LINEAR_EXACT) (8.8 fixed point) + 32 px REFLECT_101 ring
//   k_cv_fast                        whole-level FAST-9 (threshold 20) + 3x3 NMS + border 31 -> dense score map
//   k_cv_blur                        GaussianBlur(7x7, 2) as sepFilter2D's float path (see oracle/cvorb.cpp)
//   k_cv_select                      row-major compaction, retainBest(2 N_l) on the FAST score, Harris response,
//                                    retainBest(N_l): std::nth_element / std::partition permutation-exact (introselect.cuh)
//   k_cv_gather                      level lists -> frame order (+ Extractor.cpp:56-57's retainBest when ties overshoot)
//   k_cv_orient_describe             IC angle, steered BRIEF, cv::KeyPoint + Frame fields
#include "brief_pattern.h"
#include "common.cuh"
#include "fast_core.cuh"
#include "frame_tail.cuh"
#include "introselect.cuh"
#include "orb_math.cuh"
#include "tma.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#define CV_RING 32
#define CV_EDGE 31
#define CV_FAST_TH 20
#define CV_TW 64
#define CV_TH 32
#define CV_WOFF 16  // a staged window starts 16 columns left of its tile: TMA boxes start on 16-byte boundaries
#define CV_PITCH 96 // bytes per staged row: 16 + 64 + halo, multiple of 16
#define CV_BOX_ROWS 40
#define CV_SEL_THREADS 512

struct CvLevel {
    int w, h, pstride, bstride;
    unsigned pyr_off;  // byte offset of level pixel (0, 0) inside a frame slot's pyramid (32 px ring around it)
    unsigned blur_off; // blurred level and FAST score map share this geometry (w x h, pitch bstride)
    int quota;         // nfeaturesPerLevel
    int cand_off, cand_cap;
    int out_off, out_cap;
    int tabx_off, taby_off;
    int tile0;         // first 64 x 32 tile of this level in the FAST / blur grids
    int vec_end, unr_end; // sepFilter2D row pass: [0, vec_end) vector body, [vec_end, unr_end) unrolled scalar, rest tail
    float scale, inv_scale;
};
struct CvCfg {
    int nlevels, nfeatures, w, h, tiles;
    int cand_total, out_total;
    float harris_scale4;
    int umax[16];
    float gk[7];
    CvLevel lv[RGBDFE_MAX_LEVELS];
};
struct HKey {
    float r;
    unsigned xy; // y << 16 | x, level coordinates
};
struct CvSel { // a selected keypoint, level coordinates
    short x, y;
    int level;
    float response;
};

struct CvOrbWorkspace {
    CvCfg cfg;
    bool configured = false;
    int nfeatures = 0, nlevels = 0;
    float scaleFactor = 1.2f;
    size_t pyr_slot = 0, blur_slot = 0;
    uint8_t *d_pyr = nullptr, *d_blur = nullptr, *d_score = nullptr;
    unsigned* d_cand = nullptr;          // [slot][cand_total]
    unsigned short* d_lr = nullptr;      // [slot][2][cand_total] stopper positions when a level does not fit shared memory
    HKey* d_hkey = nullptr;              // [slot][cand_total] Harris-stage list (also the test tap)
    int* d_hcount = nullptr;             // [slot][levels]
    CvSel* d_out = nullptr;              // [slot][out_total]
    int* d_out_count = nullptr;          // [slot][levels], -1 = capacity overflow
    int* d_final = nullptr;              // [slot][out_total] indices into d_out in output order
    int* d_final_count = nullptr;        // [slot]
    HKey* d_gscratch = nullptr;          // [slot][out_total] k_cv_gather's rare retainBest path
    // resize tables over padded coordinates (see k_cv_pyr_level): columns (source offset, c0 | c1 << 16), row groups
    int2* d_xpad = nullptr;
    PyrGroup* d_group = nullptr;
    int xpad_off[RGBDFE_MAX_LEVELS] = {}, group_off[RGBDFE_MAX_LEVELS] = {};
    int pyr_nsrc = 0;
    int tab_cap = 0;
    int slots = 0;
    signed char* d_pattern = nullptr;
    CUtensorMap* d_maps = nullptr;       // per level: the padded pyramid level of every frame slot as a (pstride, h + 64, slots) u8 tensor, box 96 x 40
};

namespace rgbdfe {

__device__ __forceinline__ int cv_reflect(int p, int len)
{
    if (p < 0)
        p = -p;
    if (p >= len)
        p = 2 * (len - 1) - p;
    return p;
}

// ------------------------------------------------------------------ pyramid ----
// Level 0: copyMakeBorder(image, BORDER_REFLECT_101) with a 32 px ring; one thread = 4 padded bytes.
__global__ void __launch_bounds__(256) k_cv_pyr_base(const __grid_constant__ CvCfg C, const __grid_constant__ ImgPtrs img, int img_stride,
    uint8_t* __restrict__ pyr, size_t pyr_slot)
{
    const CvLevel& L = C.lv[0];
    const int slot = blockIdx.z;
    const int wq = blockIdx.x * blockDim.x + threadIdx.x; // word within the padded row
    const int py = blockIdx.y * blockDim.y + threadIdx.y - CV_RING;
    if (4 * wq >= L.pstride || py >= L.h + CV_RING)
        return;
    const uint8_t* src = static_cast<const uint8_t*>(img.p[slot]) + (size_t)cv_reflect(py, L.h) * img_stride;
    unsigned v = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int px = 4 * wq + j - CV_RING;
        if (px < L.w + CV_RING)
            v |= (unsigned)src[cv_reflect(px, L.w)] << (8 * j);
    }
    *reinterpret_cast<unsigned*>(pyr + slot * pyr_slot + L.pyr_off + (ptrdiff_t)py * L.pstride - CV_RING + 4 * wq) = v;
}

// Level l >= 1: resize(prev, INTER_LINEAR_EXACT) + copyMakeBorder(REFLECT_101 | ISOLATED).  resize_bitExact: horizontal
// blend in ufixedpoint16 (8 fractional bits), vertical blend of two such rows in ufixedpoint32, (v + 2^15) >> 16.
// Same organisation as k_pyramid_resize_direct (orb_image.cu): tables over PADDED coordinates built on the host (ring
// reflection and pitch padding resolved there; indices left of min / right of max are coefficient pairs (256, 0) /
// (0, 256) on a valid source pair), one thread = 4 padded pixels x PYR_ROWS rows, every distinct source row of the row
// group blended horizontally once, the 16-bit results parked in the thread's own shared-memory column.
struct CvPyr {
    int ph, pstride;      // padded rows and row pitch of this level
    unsigned dst_off;     // byte offset of padded (row 0, column 0) in a frame slot
    unsigned src_off;     // byte offset of level pixel (0, 0) of level l - 1
    int xpad_off, group_off;
};
template <int NSRC>
__global__ void __launch_bounds__(256) k_cv_pyr_level(const __grid_constant__ CvPyr V, uint8_t* __restrict__ pyr, size_t pyr_slot,
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
    const uint8_t* src = slot_base + V.src_off;
    const int4* xp = reinterpret_cast<const int4*>(xpad + V.xpad_off + px0);
    const int4 e0 = xp[0], e1 = xp[1];
    const int4* gp = reinterpret_cast<const int4*>(groups + V.group_off + g);
    const int4 o0 = gp[0], o1 = gp[1], bt = gp[2], m = gp[3];
    const int off[8] = { o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w };
    const int sx[4] = { e0.x, e0.z, e1.x, e1.z };
    const unsigned cx[4] = { (unsigned)e0.y, (unsigned)e0.w, (unsigned)e1.y, (unsigned)e1.w };
#pragma unroll
    for (int s = 0; s < NSRC; s++) {
        const uint8_t* row = src + off[s];
        unsigned t[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint8_t* q = row + sx[k];
            t[k] = q[0] * (cx[k] & 0xffffu) + q[1] * (cx[k] >> 16); // <= 255 * 256
        }
        s_t[s][tid] = make_uint2(t[1] * 65536u + t[0], t[3] * 65536u + t[2]);
    }
    const unsigned beta[4] = { (unsigned)bt.x, (unsigned)bt.y, (unsigned)bt.z, (unsigned)bt.w };
    uint8_t* dst = slot_base + V.dst_off + (size_t)py0 * V.pstride + px0;
#pragma unroll
    for (int j = 0; j < PYR_ROWS; j++) {
        if (py0 + j < V.ph) {
            const unsigned sel = (unsigned)m.x >> (8 * j);
            const uint2 a = s_t[sel & 15u][tid], b = s_t[(sel >> 4) & 15u][tid];
            const unsigned c0 = beta[j] & 0xffffu, c1 = beta[j] >> 16;
            const unsigned v0 = ((a.x & 0xffffu) * c0 + (b.x & 0xffffu) * c1 + 32768u) >> 16;
            const unsigned v1 = ((a.x >> 16) * c0 + (b.x >> 16) * c1 + 32768u) >> 16;
            const unsigned v2 = ((a.y & 0xffffu) * c0 + (b.y & 0xffffu) * c1 + 32768u) >> 16;
            const unsigned v3 = ((a.y >> 16) * c0 + (b.y >> 16) * c1 + 32768u) >> 16;
            *reinterpret_cast<unsigned*>(dst + (size_t)j * V.pstride) = ((v3 * 256u + v2) * 256u + v1) * 256u + v0; // bytes past the padded width are pitch padding
        }
    }
}

__device__ __forceinline__ int cv_find_level(const CvCfg& C, int tile)
{
    int level = 0;
    while (level + 1 < C.nlevels && tile >= C.lv[level + 1].tile0)
        level++;
    return level;
}

// ------------------------------------------------------------------ FAST ----
// cv::FAST(threshold 20, nonmaxSuppression) over the whole level + KeyPointsFilter::runByImageBorder(31): the tile
// computes the corner strength S on its 66 x 34 neighbourhood, applies the 3x3 non-max rule (a neighbour that is not
// a corner scores 0) and writes response = S - 1 of the survivors into the level's score map, 0 elsewhere.
__global__ void __launch_bounds__(256) k_cv_fast(const __grid_constant__ CvCfg C, const CUtensorMap* __restrict__ maps, uint8_t* __restrict__ score,
    size_t score_slot)
{
    const int level = cv_find_level(C, blockIdx.x);
    const CvLevel& L = C.lv[level];
    const int t = blockIdx.x - L.tile0;
    const int tiles_x = (L.w + CV_TW - 1) / CV_TW;
    const int ty = t / tiles_x, x0 = (t - ty * tiles_x) * CV_TW, y0 = ty * CV_TH;
    const int slot = blockIdx.y, tid = threadIdx.x;
    constexpr int PITCH = CV_PITCH, ROWS = CV_BOX_ROWS;
    __shared__ __align__(128) uint8_t s_img[ROWS][PITCH];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(16) uint8_t s_S[ROWS][PITCH];
    __shared__ __align__(16) uint8_t s_out[CV_TH][CV_TW];
    __shared__ unsigned short s_cand[(CV_TH + 2) * (CV_TW + 4)]; // at most the 34 x 66 kept region + halo
    __shared__ int s_ncand;
    uint8_t* out = score + slot * score_slot + L.blur_off;
    // the part of the tile that survives runByImageBorder(31); strengths are needed there and on its 1 px halo only
    const int ex0 = max(x0, CV_EDGE), ex1 = min(x0 + CV_TW, L.w - CV_EDGE), ey0 = max(y0, CV_EDGE), ey1 = min(y0 + CV_TH, L.h - CV_EDGE);
    if (ex0 >= ex1 || ey0 >= ey1) { // a border tile: all zeros
        for (int i = tid; i < CV_TH * (CV_TW / 4); i += 256) {
            const int rr = i >> 4, q = i & 15;
            if (y0 + rr < L.h)
                *reinterpret_cast<unsigned*>(out + (size_t)(y0 + rr) * L.bstride + x0 + 4 * q) = 0u;
        }
        return;
    }
    // staged window: level rows y0-4 .. y0+35, columns x0-16 .. x0+79, brought in by the TMA unit (box 96 x 40; whatever
    // lies beyond the padded level arrives as zeros and is outside the detection region); the threads clear the strength
    // map and the output tile meanwhile
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_expect_tx(&s_bar, PITCH * ROWS);
        tma_load_3d(&s_img[0][0], &maps[level], x0 - CV_WOFF + CV_RING, y0 - 4 + CV_RING, slot, &s_bar);
    }
    for (int i = tid; i < ROWS * (PITCH / 4); i += 256)
        reinterpret_cast<unsigned*>(&s_S[0][0])[i] = 0u;
    for (int i = tid; i < CV_TH * (CV_TW / 4); i += 256)
        reinterpret_cast<unsigned*>(&s_out[0][0])[i] = 0u;
    if (tid == 0)
        s_ncand = 0;
    __syncthreads();
    mbar_wait(&s_bar, 0);
    // Phase 1: the 4-opposite-pair necessary test (fast_candidate2, four pixels per thread in 16-bit lanes) on the kept
    // region + 1 px halo (which lies inside cv::FAST's detection region [3, w-3) x [3, h-3)); survivors are compacted
    // with ballots.  Window column c is level column x0 - 16 + c, window row r is level row y0 - 4 + r.
    {
        constexpr int RW = PITCH / 4;
        const int r_lo = ey0 - 1 - (y0 - 4), nrows = ey1 - ey0 + 2;
        const int c_lo = ex0 - 1 - (x0 - CV_WOFF), c_hi = ex1 - (x0 - CV_WOFF); // first / last needed window column
        const int g_lo = c_lo >> 2, ngr = (c_hi >> 2) - g_lo + 1;  // <= 18 of the 24 word groups of a row
        const unsigned inv_ngr = 0xffffffffu / (unsigned)ngr + 1u;
        const int NG = nrows * ngr;
        const unsigned c1 = 0x7fff7fffu - ((unsigned)CV_FAST_TH | (unsigned)CV_FAST_TH << 16);
        const unsigned lt = (1u << (tid & 31)) - 1;
        for (int i0 = (tid & ~31); i0 < NG; i0 += 256) { // warp-uniform trip count
            const int i = i0 + (tid & 31);
            unsigned rA = 0, rB = 0;
            int r = 0, g = 0;
            if (i < NG) {
                const int rr = (int)__umulhi((unsigned)i, inv_ngr);
                r = r_lo + rr;
                g = g_lo + (i - rr * ngr);
                const unsigned* w = reinterpret_cast<const unsigned*>(&s_img[r][0]) + g;
                const unsigned cm = w[-1], cc = w[0], cp = w[1];
                const unsigned up3 = w[-3 * RW], dn3 = w[3 * RW];
                const unsigned um = w[-2 * RW - 1], uc = w[-2 * RW], upn = w[-2 * RW + 1];
                const unsigned dm = w[2 * RW - 1], dc = w[2 * RW], dpn = w[2 * RW + 1];
                const unsigned u4 = __funnelshift_r(cc, cp, 24), u12 = __funnelshift_r(cm, cc, 8);
                const unsigned u2 = __funnelshift_r(dc, dpn, 16), u14 = __funnelshift_r(dm, dc, 16);
                const unsigned u6 = __funnelshift_r(uc, upn, 16), u10 = __funnelshift_r(um, uc, 16);
                rA = fast_candidate2(lanes_lo(cc), lanes_lo(dn3), lanes_lo(up3), lanes_lo(u4), lanes_lo(u12), lanes_lo(u2), lanes_lo(u10),
                    lanes_lo(u6), lanes_lo(u14), c1);
                rB = fast_candidate2(lanes_hi(cc), lanes_hi(dn3), lanes_hi(up3), lanes_hi(u4), lanes_hi(u12), lanes_hi(u2), lanes_hi(u10),
                    lanes_hi(u6), lanes_hi(u14), c1);
            }
            const int cb = 4 * g;
            const bool k0 = (rA & 0x8000u) && cb >= c_lo && cb <= c_hi, k1 = (rA >> 31) && cb + 1 >= c_lo && cb + 1 <= c_hi;
            const bool k2 = (rB & 0x8000u) && cb + 2 >= c_lo && cb + 2 <= c_hi, k3 = (rB >> 31) && cb + 3 >= c_lo && cb + 3 <= c_hi;
            const unsigned b0 = __ballot_sync(0xffffffffu, k0), b1 = __ballot_sync(0xffffffffu, k1);
            const unsigned b2 = __ballot_sync(0xffffffffu, k2), b3 = __ballot_sync(0xffffffffu, k3);
            const int n0 = __popc(b0), n1 = __popc(b1), n2 = __popc(b2), n3 = __popc(b3);
            if (b0 | b1 | b2 | b3) {
                int base = 0;
                if ((tid & 31) == 0)
                    base = atomicAdd(&s_ncand, n0 + n1 + n2 + n3);
                base = __shfl_sync(0xffffffffu, base, 0);
                const int key = r << 8 | cb;
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
    }
    __syncthreads();
    // Phase 2: corner strength of the candidates (dense warps)
    const int ncand = s_ncand;
    for (int i = tid; i < ncand; i += 256) {
        const int r = s_cand[i] >> 8, c = s_cand[i] & 255;
        const int S = fast_strength(&s_img[r][c], PITCH, CV_FAST_TH);
        s_S[r][c] = (uint8_t)min(S, 255);
    }
    __syncthreads();
    // Phase 3: 3x3 non-max suppression of the corners inside the kept region (a neighbour that is not a corner holds 0,
    // a corner's S is >= 21, and score = S - 1 on both sides of the comparison)
    for (int i = tid; i < ncand; i += 256) {
        const int r = s_cand[i] >> 8, c = s_cand[i] & 255;
        const int x = x0 - CV_WOFF + c, y = y0 - 4 + r;
        const int S = s_S[r][c];
        if (S == 0 || x < ex0 || x >= ex1 || y < ey0 || y >= ey1)
            continue;
        const int m = max(max(max(s_S[r - 1][c - 1], s_S[r - 1][c]), max(s_S[r - 1][c + 1], s_S[r][c - 1])),
            max(max(s_S[r][c + 1], s_S[r + 1][c - 1]), max(s_S[r + 1][c], s_S[r + 1][c + 1])));
        if (S > m)
            s_out[r - 4][c - CV_WOFF] = (uint8_t)(S - 1);
    }
    __syncthreads();
    for (int i = tid; i < CV_TH * (CV_TW / 4); i += 256) {
        const int rr = i >> 4, q = i & 15;
        if (y0 + rr < L.h)
            *reinterpret_cast<unsigned*>(out + (size_t)(y0 + rr) * L.bstride + x0 + 4 * q) = *reinterpret_cast<const unsigned*>(&s_out[rr][4 * q]);
    }
}

// ------------------------------------------------------------------ blur ----
// GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) on a submatrix = sepFilter2D with the float kernel: row pass
// left to right (FMA in the 32-pixel vector body and in the last w % 4 pixels, separate multiply / add in the
// 4-way unrolled part between them), column pass centre tap + symmetric pairs with FMA, cvRound, saturate.
__global__ void __launch_bounds__(256) k_cv_blur(const __grid_constant__ CvCfg C, const CUtensorMap* __restrict__ maps, uint8_t* __restrict__ blur,
    size_t blur_slot)
{
    const int level = cv_find_level(C, blockIdx.x);
    const CvLevel& L = C.lv[level];
    const int t = blockIdx.x - L.tile0;
    const int tiles_x = (L.w + CV_TW - 1) / CV_TW;
    const int ty = t / tiles_x, x0 = (t - ty * tiles_x) * CV_TW, y0 = ty * CV_TH;
    const int slot = blockIdx.y, tid = threadIdx.x;
    constexpr int PITCH = CV_PITCH, ROWS = CV_TH + 6;
    __shared__ __align__(128) uint8_t s_img[CV_BOX_ROWS][PITCH];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(16) float s_h[ROWS][CV_TW];
    // staged window: level rows y0-3 .. y0+36, columns x0-16 .. x0+79 (TMA box 96 x 40, see k_cv_fast)
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_expect_tx(&s_bar, PITCH * CV_BOX_ROWS);
        tma_load_3d(&s_img[0][0], &maps[level], x0 - CV_WOFF + CV_RING, y0 - 3 + CV_RING, slot, &s_bar);
    }
    __syncthreads();
    mbar_wait(&s_bar, 0);
    // row pass: 4 adjacent outputs per thread from 3 staged words; a group of 4 never straddles vec_end / unr_end
    for (int i = tid; i < ROWS * (CV_TW / 4); i += 256) {
        const int r = i >> 4, q = i & 15;
        const unsigned* row = reinterpret_cast<const unsigned*>(&s_img[r][0]);
        const unsigned w0 = row[q + 3], w1 = row[q + 4], w2 = row[q + 5]; // level columns x0 + 4q - 4 .. x0 + 4q + 7
        float p[12];
#pragma unroll
        for (int j = 0; j < 4; j++) { // byte -> float on the ALU: 0x4B0000bb is 2^23 + b (keeps the conversion off the XU pipe)
            p[j] = __fsub_rn(__uint_as_float(__byte_perm(w0, 0x4B000000u, 0x7650 + j)), 8388608.0f);
            p[4 + j] = __fsub_rn(__uint_as_float(__byte_perm(w1, 0x4B000000u, 0x7650 + j)), 8388608.0f);
            p[8 + j] = __fsub_rn(__uint_as_float(__byte_perm(w2, 0x4B000000u, 0x7650 + j)), 8388608.0f);
        }
        const int x = x0 + 4 * q;
        float acc[4];
        if (x < L.vec_end || x >= L.unr_end) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                acc[j] = __fmul_rn(C.gk[0], p[j + 1]); // staged column c + 1 is level column x - 3
#pragma unroll
                for (int t = 1; t < 7; t++)
                    acc[j] = __fmaf_rn(C.gk[t], p[j + 1 + t], acc[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                acc[j] = __fmul_rn(C.gk[0], p[j + 1]);
#pragma unroll
                for (int t = 1; t < 7; t++)
                    acc[j] = __fadd_rn(acc[j], __fmul_rn(C.gk[t], p[j + 1 + t]));
            }
        }
        *reinterpret_cast<float4*>(&s_h[r][4 * q]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
    __syncthreads();
    // column pass: one thread = one column x 8 output rows (14 row sums held in registers); results staged as bytes
    uint8_t(*s_out)[CV_TW] = reinterpret_cast<uint8_t(*)[CV_TW]>(&s_img[0][0]); // the staged image is dead
    {
        const int c = tid & 63, band = tid >> 6;
        float v[14];
#pragma unroll
        for (int k = 0; k < 14; k++)
            v[k] = s_h[band * 8 + k][c];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            float acc = __fmul_rn(C.gk[3], v[k + 3]);
#pragma unroll
            for (int t = 1; t <= 3; t++)
                acc = __fmaf_rn(C.gk[3 + t], __fadd_rn(v[k + 3 + t], v[k + 3 - t]), acc);
            // cvRound = round-half-even: adding 1.5 * 2^23 leaves the rounded integer in the low mantissa bits
            const int iv = __float_as_int(__fadd_rn(acc, 12582912.0f)) - 0x4B400000;
            s_out[band * 8 + k][c] = (uint8_t)min(max(iv, 0), 255);
        }
    }
    __syncthreads();
    uint8_t* out = blur + slot * blur_slot + L.blur_off;
    for (int i = tid; i < CV_TH * (CV_TW / 4); i += 256) {
        const int rr = i >> 4, q = i & 15;
        const int y = y0 + rr;
        if (y < L.h)
            *reinterpret_cast<unsigned*>(out + (size_t)y * L.bstride + x0 + 4 * q) = *reinterpret_cast<const unsigned*>(&s_out[rr][4 * q]);
    }
}

// ------------------------------------------------------------------ selection ----
struct FastGreater { // KeypointResponseGreater on packed candidates (response in the top byte)
    __device__ bool operator()(unsigned a, unsigned b) const { return (a >> 24) > (b >> 24); }
};
struct HarrisGreater {
    __device__ bool operator()(const HKey& a, const HKey& b) const { return a.r > b.r; }
};

// cv::KeyPointsFilter::retainBest by the whole CTA (every thread calls it); returns the new size.
template <class T, class C, class P>
__device__ int cta_retain_best(T* a, int n, int n_points, unsigned short* Lpos, unsigned short* Rpos, int* s_w, C comp, P make_pred)
{
    if (n_points < 0 || n <= n_points)
        return n;
    if (n_points == 0)
        return 0;
    cta_nth_element(a, n_points - 1, n, Lpos, Rpos, s_w, comp);
    const T amb = a[n_points - 1];
    return cta_partition(a, n_points, n, Lpos, Rpos, s_w, make_pred(amb));
}

__device__ __forceinline__ int cv_nz(unsigned v) { return __popc(__vcmpne4(v, 0u)) >> 3; } // non-zero bytes of a word
__device__ __forceinline__ int cv_nz4(const uint4& v) { return cv_nz(v.x) + cv_nz(v.y) + cv_nz(v.z) + cv_nz(v.w); }

// Row-major compaction of a level's score map.  A lane owns UPL consecutive 16-byte units of a row (so one warp scan
// per row keeps the order) and a warp keeps R = 8 / UPL rows in flight to overlap the memory latency.
// Pass 1 leaves the per-row counts in s_rowoff; pass 2 (after the scan) writes the packed candidates.
template <int UPL>
__device__ void cv_count_rows(const uint8_t* sc, int bstride, int h, int units, int* s_rowoff, int warp, int nwarps, int lane)
{
    constexpr int R = 8 / UPL;
    const int ua = lane * UPL;
    for (int yb = warp * R; yb < h; yb += nwarps * R) {
        uint4 seg[R][UPL];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int y = yb + r;
            const bool in = y >= CV_EDGE && y < h - CV_EDGE; // rows outside the border filter hold zeros
#pragma unroll
            for (int k = 0; k < UPL; k++)
                seg[r][k] = (in && ua + k < units) ? reinterpret_cast<const uint4*>(sc + (size_t)y * bstride)[ua + k] : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            int cnt = 0;
#pragma unroll
            for (int k = 0; k < UPL; k++)
                cnt += cv_nz4(seg[r][k]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            if (lane == 0 && yb + r < h)
                s_rowoff[yb + r] = cnt;
        }
    }
}

template <int UPL>
__device__ void cv_emit_rows(const uint8_t* sc, int bstride, int h, int units, const int* s_rowoff, unsigned* keys, int warp, int nwarps, int lane)
{
    constexpr int R = 8 / UPL;
    const int ua = lane * UPL;
    for (int yb = warp * R; yb < h; yb += nwarps * R) {
        uint4 seg[R][UPL];
        const int last = min(yb + R, h);
        if (s_rowoff[last] == s_rowoff[yb])
            continue;
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int y = yb + r;
            const bool in = y < h && s_rowoff[min(y + 1, h)] != s_rowoff[min(y, h)];
#pragma unroll
            for (int k = 0; k < UPL; k++)
                seg[r][k] = (in && ua + k < units) ? reinterpret_cast<const uint4*>(sc + (size_t)y * bstride)[ua + k] : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int y = yb + r;
            if (y >= h)
                break;
            int c = 0;
#pragma unroll
            for (int k = 0; k < UPL; k++)
                c += cv_nz4(seg[r][k]);
            int inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o)
                    inc += u;
            }
            int w = s_rowoff[y] + inc - c;
            if (c) {
#pragma unroll
                for (int k = 0; k < UPL; k++) {
                    const unsigned vv[4] = { seg[r][k].x, seg[r][k].y, seg[r][k].z, seg[r][k].w };
#pragma unroll
                    for (int m = 0; m < 4; m++) {
                        unsigned mk = __vcmpne4(vv[m], 0u); // 0xff per surviving corner (at most 2 per word: NMS)
                        while (mk) {
                            const int j = (__ffs(mk) - 1) >> 3;
                            keys[w++] = ((vv[m] >> (8 * j)) & 0xffu) << 24 | (unsigned)y << 12 | (unsigned)(16 * (ua + k) + 4 * m + j);
                            mk &= ~(0xffu << (8 * j));
                        }
                    }
                }
            }
        }
    }
}

// One CTA per (level, frame).  Dynamic shared memory: keys[kcap] u32 | Lpos[kcap] u16 | Rpos[kcap] u16 | rowoff[h + 1];
// kcap = FAST candidates of one level handled in shared memory (sized from the image so that several CTAs share an SM);
// a level with more candidates works in the global scratch instead.  The Harris stage reuses the key area (kcap / 2 entries).
__global__ void __launch_bounds__(CV_SEL_THREADS, 2) k_cv_select(const __grid_constant__ CvCfg C, const uint8_t* __restrict__ pyr, size_t pyr_slot,
    const uint8_t* __restrict__ score, size_t score_slot, unsigned* __restrict__ cand_all, unsigned short* __restrict__ lr_all,
    HKey* __restrict__ hkey_all, int* __restrict__ hcount, CvSel* __restrict__ out_all, int* __restrict__ out_count, int kcap)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned* s_keys = reinterpret_cast<unsigned*>(smem);
    unsigned short* s_L = reinterpret_cast<unsigned short*>(smem + (size_t)kcap * 4);
    unsigned short* s_R = s_L + kcap;
    int* s_rowoff = reinterpret_cast<int*>(smem + (size_t)kcap * 8);
    __shared__ int s_n, s_carry;
    __shared__ int s_w[68];
    const int level = blockIdx.x, slot = blockIdx.y;
    const CvLevel& L = C.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = CV_SEL_THREADS / 32;
    const uint8_t* sc = score + slot * score_slot + L.blur_off;

    // ---- row-major compaction of the score map (cv::FAST emits keypoints row by row) ----
    const int units = (L.w + 15) >> 4, upl = (units + 31) >> 5; // 16-byte units per row / per lane (<= 8 for w <= 4095)
    if (upl <= 1)
        cv_count_rows<1>(sc, L.bstride, L.h, units, s_rowoff, warp, nwarps, lane);
    else if (upl <= 2)
        cv_count_rows<2>(sc, L.bstride, L.h, units, s_rowoff, warp, nwarps, lane);
    else if (upl <= 4)
        cv_count_rows<4>(sc, L.bstride, L.h, units, s_rowoff, warp, nwarps, lane);
    else
        cv_count_rows<8>(sc, L.bstride, L.h, units, s_rowoff, warp, nwarps, lane);
    if (tid == 0)
        s_carry = 0;
    __syncthreads();
    if (warp == 0) { // exclusive scan over the rows
        int carry = 0;
        for (int y0 = 0; y0 < L.h; y0 += 32) {
            const int y = y0 + lane;
            const int v = y < L.h ? s_rowoff[y] : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o)
                    inc += u;
            }
            if (y < L.h)
                s_rowoff[y] = carry + inc - v;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) {
            s_rowoff[L.h] = carry;
            s_n = carry;
        }
    }
    __syncthreads();
    int n = s_n;
    const bool overflow = n > L.cand_cap || n > 65535;
    if (overflow)
        n = 0; // reported through out_count = -1
    unsigned* g_keys = cand_all + (size_t)slot * C.cand_total + L.cand_off;
    const bool in_smem = n <= kcap;
    unsigned* keys = in_smem ? s_keys : g_keys;
    unsigned short* Lpos = in_smem ? s_L : lr_all + ((size_t)slot * 2) * C.cand_total + L.cand_off;
    unsigned short* Rpos = in_smem ? s_R : lr_all + ((size_t)slot * 2 + 1) * C.cand_total + L.cand_off;
    if (n > 0) {
        if (upl <= 1)
            cv_emit_rows<1>(sc, L.bstride, L.h, units, s_rowoff, keys, warp, nwarps, lane);
        else if (upl <= 2)
            cv_emit_rows<2>(sc, L.bstride, L.h, units, s_rowoff, keys, warp, nwarps, lane);
        else if (upl <= 4)
            cv_emit_rows<4>(sc, L.bstride, L.h, units, s_rowoff, keys, warp, nwarps, lane);
        else
            cv_emit_rows<8>(sc, L.bstride, L.h, units, s_rowoff, keys, warp, nwarps, lane);
    }
    __syncthreads();

    // ---- retainBest(2 * N_l) on the FAST score ----
    const int n1 = cta_retain_best(keys, n, 2 * L.quota, Lpos, Rpos, s_w, FastGreater(),
        [](unsigned amb) { return [amb](unsigned k) { return (k >> 24) >= (amb >> 24); }; });

    // ---- HarrisResponses(blockSize 7, k 0.04) ----
    HKey* g_h = hkey_all + (size_t)slot * C.cand_total + L.cand_off;
    const uint8_t* img = pyr + slot * pyr_slot + L.pyr_off;
    const int step = L.pstride;
    // one warp per keypoint, four keypoints in flight: lane p (and p + 32) takes pixel p of the 7 x 7 block, the integer
    // sums are reduced with shuffles
    for (int i0 = warp * 4; i0 < n1; i0 += nwarps * 4) {
        int a[4], b[4], c[4];
        unsigned kk[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            a[q] = b[q] = c[q] = 0;
            kk[q] = keys[min(i0 + q, n1 - 1)];
            const int x0 = kk[q] & 0xfff, y0 = (kk[q] >> 12) & 0xfff;
            const uint8_t* ptr0 = img + (ptrdiff_t)(y0 - 3) * step + x0 - 3;
#pragma unroll
            for (int rnd = 0; rnd < 2; rnd++) {
                const int p = lane + 32 * rnd;
                if (p < 49) {
                    const int r = p / 7, j = p - 7 * r;
                    const uint8_t* ptr = ptr0 + (ptrdiff_t)r * step + j;
                    const int m0 = ptr[-step - 1], m1 = ptr[-step], m2 = ptr[-step + 1];
                    const int c0 = ptr[-1], c2 = ptr[1];
                    const int p0 = ptr[step - 1], p1 = ptr[step], p2 = ptr[step + 1];
                    const int Ix = (c2 - c0) * 2 + (m2 - m0) + (p2 - p0);
                    const int Iy = (p1 - m1) * 2 + (p0 - m0) + (p2 - m2);
                    a[q] += Ix * Ix;
                    b[q] += Iy * Iy;
                    c[q] += Ix * Iy;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a[q] += __shfl_xor_sync(0xffffffffu, a[q], o);
                b[q] += __shfl_xor_sync(0xffffffffu, b[q], o);
                c[q] += __shfl_xor_sync(0xffffffffu, c[q], o);
            }
            if (lane == q && i0 + q < n1) {
                const float fa = (float)a[q], fb = (float)b[q], fc = (float)c[q];
                const float sum = __fadd_rn(fa, fb);
                const float resp
                    = __fmul_rn(__fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)), __fmul_rn(__fmul_rn(0.04f, sum), sum)), C.harris_scale4);
                HKey hh;
                hh.r = resp;
                hh.xy = ((kk[q] >> 12) & 0xfff) << 16 | (kk[q] & 0xfff);
                g_h[i0 + q] = hh;
            }
        }
    }
    if (tid == 0)
        hcount[slot * RGBDFE_MAX_LEVELS + level] = n1;
    __syncthreads();
    // ---- retainBest(N_l) on the Harris response ----
    const bool h_in_smem = n1 <= kcap / 2;
    HKey* hk = h_in_smem ? reinterpret_cast<HKey*>(smem) : g_h; // the FAST keys are dead: their shared memory is reused
    if (h_in_smem)
        for (int i = tid; i < n1; i += CV_SEL_THREADS)
            hk[i] = g_h[i];
    __syncthreads();
    // stopper lists: the Harris list never exceeds the candidate capacity, so the global scratch always fits
    unsigned short* L2 = h_in_smem ? s_L : lr_all + ((size_t)slot * 2) * C.cand_total + L.cand_off;
    unsigned short* R2 = h_in_smem ? s_R : lr_all + ((size_t)slot * 2 + 1) * C.cand_total + L.cand_off;
    const int n2 = cta_retain_best(hk, n1, L.quota, L2, R2, s_w, HarrisGreater(), [](HKey amb) { return [amb](const HKey& k) { return k.r >= amb.r; }; });
    CvSel* out = out_all + (size_t)slot * C.out_total + L.out_off;
    for (int i = tid; i < min(n2, L.out_cap); i += CV_SEL_THREADS) {
        CvSel s;
        s.x = (short)(hk[i].xy & 0xffff);
        s.y = (short)(hk[i].xy >> 16);
        s.level = level;
        s.response = hk[i].r;
        out[i] = s;
    }
    if (tid == 0)
        out_count[slot * RGBDFE_MAX_LEVELS + level] = (overflow || n2 > L.out_cap) ? -1 : n2;
}

// One CTA per frame: output order = levels in sequence.  If ties pushed the total above nfeatures, Extractor.cpp:56-57
// runs retainBest(nfeatures) over the whole list and cv::ORB's compute() then regroups the survivors by level
// (stable); that rare path runs on one thread.
__global__ void __launch_bounds__(256) k_cv_gather(const __grid_constant__ CvCfg C, const CvSel* __restrict__ out_all, const int* __restrict__ out_count,
    int* __restrict__ final_all, int* __restrict__ final_count, HKey* __restrict__ scratch_all)
{
    const int slot = blockIdx.x, tid = threadIdx.x;
    __shared__ int s_off[RGBDFE_MAX_LEVELS + 1];
    __shared__ int s_bad;
    if (tid == 0) {
        int acc = 0, bad = 0;
        for (int l = 0; l < C.nlevels; l++) {
            s_off[l] = acc;
            const int n = out_count[slot * RGBDFE_MAX_LEVELS + l];
            if (n < 0)
                bad = 1;
            acc += max(n, 0);
        }
        s_off[C.nlevels] = acc;
        s_bad = bad;
    }
    __syncthreads();
    int* fin = final_all + (size_t)slot * C.out_total;
    const int total = s_off[C.nlevels];
    if (s_bad) {
        if (tid == 0)
            final_count[slot] = -1;
        return;
    }
    if (total <= C.nfeatures) {
        for (int l = 0; l < C.nlevels; l++)
            for (int i = tid; i < s_off[l + 1] - s_off[l]; i += 256)
                fin[s_off[l] + i] = C.lv[l].out_off + i;
        if (tid == 0)
            final_count[slot] = total;
        return;
    }
    if (tid == 0) {
        HKey* g = scratch_all + (size_t)slot * C.out_total;
        const CvSel* out = out_all + (size_t)slot * C.out_total;
        int n = 0;
        for (int l = 0; l < C.nlevels; l++)
            for (int i = 0; i < s_off[l + 1] - s_off[l]; i++) {
                g[n].r = out[C.lv[l].out_off + i].response;
                g[n].xy = (unsigned)(C.lv[l].out_off + i);
                n++;
            }
        struct G {
            __device__ bool operator()(const HKey& a, const HKey& b) const { return a.r > b.r; }
        };
        std_nth_element(g, C.nfeatures - 1, n, G());
        const float amb = g[C.nfeatures - 1].r;
        const int m = std_partition(g, C.nfeatures, n, [amb](const HKey& k) { return k.r >= amb; });
        int w = 0;
        for (int l = 0; l < C.nlevels; l++)
            for (int i = 0; i < m; i++)
                if (out[g[i].xy].level == l)
                    fin[w++] = (int)g[i].xy;
        final_count[slot] = m;
    }
}

// ------------------------------------------------------------------ orientation + descriptor ----
__global__ void __launch_bounds__(256) k_cv_orient_describe(const __grid_constant__ CvCfg C, const uint8_t* __restrict__ pyr, size_t pyr_slot,
    const uint8_t* __restrict__ blur, size_t blur_slot, const CvSel* __restrict__ out_all, const int* __restrict__ final_all,
    const int* __restrict__ final_count, const signed char* __restrict__ pattern, const FrameBatch frames, const __grid_constant__ ImgPtrs depth,
    int depth_stride, int unproject, const __grid_constant__ FramePrep prep)
{
    __shared__ __align__(16) signed char s_pattern[1024];
    reinterpret_cast<int*>(s_pattern)[threadIdx.x] = reinterpret_cast<const int*>(pattern)[threadIdx.x];
    __syncthreads();
    const int slot = blockIdx.y, lane = threadIdx.x & 31;
    const int gi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const rgbdfe_frame F = frames.f[slot];
    const int total = final_count[slot];
    if (gi == 0 && lane == 0)
        *F.n = total < 0 ? -1 : min(total, F.cap);
    if (gi >= total || gi >= F.cap)
        return;
    const CvSel s = out_all[(size_t)slot * C.out_total + final_all[(size_t)slot * C.out_total + gi]];
    const CvLevel& L = C.lv[s.level];
    // ICAngles: integer moments over the 31 px disc of the un-blurred level
    const uint8_t* center = pyr + slot * pyr_slot + L.pyr_off + (ptrdiff_t)s.y * L.pstride + s.x;
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int u = lane - 15, au = abs(u);
        m10 = u * center[u];
        for (int v = 1; v <= 15; v++)
            if (au <= C.umax[v]) {
                const int vp = center[u + v * L.pstride], vm = center[u - v * L.pstride];
                m10 += u * (vp + vm);
                m01 += v * (vp - vm);
            }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);
    const float ptx = __fmul_rn((float)s.x, L.scale), pty = __fmul_rn((float)s.y, L.scale); // allKeypoints[i].pt *= scale
    // computeOrbDescriptors: centre = cvRound(pt * (1 / layerScale)) on the blurred level
    const int cx = __float2int_rn(__fmul_rn(ptx, L.inv_scale)), cy = __float2int_rn(__fmul_rn(pty, L.inv_scale));
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    sincosf_glibc(__fmul_rn(angle, factorPI), &b, &a);
    const uint8_t* bc = blur + slot * blur_slot + L.blur_off + (size_t)cy * L.bstride + cx;
    signed char pat[32];
    *reinterpret_cast<uint4*>(pat) = *reinterpret_cast<const uint4*>(s_pattern + lane * 32);
    *reinterpret_cast<uint4*>(pat + 16) = *reinterpret_cast<const uint4*>(s_pattern + lane * 32 + 16);
    unsigned val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float x0 = (float)pat[k * 4], y0 = (float)pat[k * 4 + 1], x1 = (float)pat[k * 4 + 2], y1 = (float)pat[k * 4 + 3];
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        const int t0 = bc[r0 * L.bstride + c0], t1 = bc[r1 * L.bstride + c1];
        val |= (unsigned)(t0 < t1) << k;
    }
    F.desc[(size_t)gi * 32 + lane] = (uint8_t)val;
    if (lane == 0) {
        rgbdfe_keypoint kp;
        kp.x = ptx;
        kp.y = pty;
        kp.size = __fmul_rn(31.0f, L.scale);
        kp.angle = angle;
        kp.response = s.response;
        kp.octave = s.level;
        kp.class_id = -1;
        store_frame_keypoint(F, gi, kp, prep, depth.p[slot], depth_stride, unproject);
    }
}

// ------------------------------------------------------------------ host ----
namespace {
template <class T>
cudaError_t dmalloc(T** p, size_t bytes)
{
    return cudaMalloc(reinterpret_cast<void**>(p), bytes);
}
void free_buffers(CvOrbWorkspace* W)
{
    cudaFree(W->d_pyr), cudaFree(W->d_blur), cudaFree(W->d_score), cudaFree(W->d_cand), cudaFree(W->d_lr), cudaFree(W->d_hkey);
    cudaFree(W->d_hcount), cudaFree(W->d_out), cudaFree(W->d_out_count), cudaFree(W->d_final), cudaFree(W->d_final_count);
    cudaFree(W->d_xpad), cudaFree(W->d_group), cudaFree(W->d_gscratch), cudaFree(W->d_maps);
    W->d_gscratch = nullptr, W->d_maps = nullptr;
    W->d_pyr = W->d_blur = W->d_score = nullptr;
    W->d_cand = nullptr, W->d_lr = nullptr, W->d_hkey = nullptr, W->d_hcount = nullptr, W->d_out = nullptr, W->d_out_count = nullptr;
    W->d_final = nullptr, W->d_final_count = nullptr, W->d_xpad = nullptr, W->d_group = nullptr;
}
int cv_round_f(float v) { return (int)lrintf(v); }
// interpolationLinear<uchar>::getCoeffs (resize.cpp), arithmetic in IEEE double (= softdouble)
void exact_table(int srcsize, int dstsize, int* ofs, unsigned short* c1)
{
    const double scale = 1.0 / ((double)dstsize / srcsize);
    for (int v = 0; v < dstsize; v++) {
        const double fval = scale * ((double)v + 0.5) - 0.5;
        const int ival = (int)std::floor(fval);
        if (ival >= 0 && srcsize > 1) {
            if (ival < srcsize - 1) {
                ofs[v] = ival;
                const double frac = fval - (double)ival;
                c1[v] = frac < 0 ? 0 : (unsigned short)std::nearbyint(frac * 256.0);
            } else { // right of max: src[last] -- as pair (last - 1, last) with coefficients (0, 256)
                ofs[v] = srcsize - 2;
                c1[v] = 256;
            }
        } else { // left of min: src[0]
            ofs[v] = 0;
            c1[v] = 0;
        }
    }
}
} // namespace

void free_cvorb_workspace(rgbdfe_ctx* c)
{
    if (!c->cvorb)
        return;
    free_buffers(c->cvorb);
    cudaFree(c->cvorb->d_pattern);
    delete c->cvorb;
    c->cvorb = nullptr;
}

int cvorb_configure(rgbdfe_ctx* c, int nfeatures, float scaleFactor, int nlevels)
{
    if (nfeatures < 1 || nfeatures > c->max_feats || nlevels < 1 || nlevels > RGBDFE_MAX_LEVELS || !(scaleFactor > 1.0f)) {
        c->set_error("cvorb_configure: nfeatures in [1, max_feats], nlevels in [1, 12], scaleFactor > 1");
        return RGBDFE_E_ARG;
    }
    if (scaleFactor != 1.2f) {
        // the reference's descriptor object is a default cv::ORB (scale 1.2f, Extractor.cpp:217): with another detector
        // scale its compute() would sample a different pyramid than detect() used
        c->set_error("cvorb_configure: the ORB/ORB pair of the reference is only consistent for scaleFactor == 1.2f");
        return RGBDFE_E_ARG;
    }
    if (!c->cvorb)
        c->cvorb = new CvOrbWorkspace();
    CvOrbWorkspace* W = c->cvorb;
    W->nfeatures = nfeatures, W->nlevels = nlevels, W->scaleFactor = scaleFactor;
    W->configured = true;
    W->cfg.w = W->cfg.h = 0; // tables are rebuilt on the next extraction
    return RGBDFE_OK;
}

static int cvorb_prepare(rgbdfe_ctx* c, int w, int h)
{
    CvOrbWorkspace* W = c->cvorb;
    if (W->cfg.w == w && W->cfg.h == h && W->slots == c->max_batch)
        return RGBDFE_OK;
    if (w > c->max_w || h > c->max_h || w > 4095 || h > 4095) {
        c->set_error("cvorb: image larger than the context (or than 4095 px)");
        return RGBDFE_E_ARG;
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    free_buffers(W);
    CvCfg& C = W->cfg;
    memset(&C, 0, sizeof(C));
    C.nlevels = W->nlevels, C.nfeatures = W->nfeatures, C.w = w, C.h = h;
    const double sf = (double)W->scaleFactor;
    // nfeaturesPerLevel (orb.cpp computeKeyPoints)
    {
        const float factor = (float)(1.0 / sf);
        float ndesired = W->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)W->nlevels));
        int sum = 0;
        for (int l = 0; l < W->nlevels - 1; l++) {
            C.lv[l].quota = cv_round_f(ndesired);
            sum += C.lv[l].quota;
            ndesired *= factor;
        }
        C.lv[W->nlevels - 1].quota = std::max(W->nfeatures - sum, 0);
    }
    {
        const int half = 15;
        int umax[17] = { 0 };
        int v, v0, vmax = (int)std::floor(half * std::sqrt(2.f) / 2 + 1), vmin = (int)std::ceil(half * std::sqrt(2.f) / 2);
        for (v = 0; v <= vmax; ++v)
            umax[v] = (int)std::nearbyint(std::sqrt((double)half * half - v * v));
        for (v = half, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1])
                ++v0;
            umax[v] = v0;
            ++v0;
        }
        for (v = 0; v < 16; v++)
            C.umax[v] = umax[v];
    }
    {
        double t[7], sum = 0; // getGaussianKernel(7, 2, CV_32F)
        for (int i = 0; i < 7; i++) {
            const double x = i - 3.0;
            t[i] = std::exp(-0.5 / 4.0 * x * x);
            sum += t[i];
        }
        sum = 1. / sum;
        for (int i = 0; i < 7; i++)
            C.gk[i] = (float)(t[i] * sum);
        const float s = 1.f / ((1 << 2) * 7 * 255.f);
        C.harris_scale4 = s * s * s * s;
    }
    size_t pyr = 0, blur = 0;
    int tiles = 0, cand = 0, out = 0, tab = 0;
    for (int l = 0; l < C.nlevels; l++) {
        CvLevel& L = C.lv[l];
        L.scale = (float)std::pow(sf, (double)l);
        L.inv_scale = 1.0f / L.scale;
        L.w = cv_round_f((float)w * L.inv_scale), L.h = cv_round_f((float)h * L.inv_scale);
        if (L.w < CV_RING + 2 || L.h < CV_RING + 2) { // the REFLECT_101 ring must fold at most once
            c->set_error("cvorb: image too small for the pyramid");
            return RGBDFE_E_ARG;
        }
        L.pstride = align_up(L.w + 2 * CV_RING, 64);
        L.bstride = align_up(L.w, 64);
        L.pyr_off = (unsigned)(pyr + (size_t)CV_RING * L.pstride + CV_RING);
        pyr += (size_t)L.pstride * (L.h + 2 * CV_RING);
        L.blur_off = (unsigned)blur;
        blur += (size_t)L.bstride * align_up(L.h, CV_TH);
        L.tile0 = tiles;
        tiles += ((L.w + CV_TW - 1) / CV_TW) * ((L.h + CV_TH - 1) / CV_TH);
        L.cand_off = cand;
        L.cand_cap = std::min(65535, std::max(4096, L.w * L.h / 8)); // 3x3 NMS leaves at most 1 corner per 4 pixels; white noise reaches 1 per 13
        cand += align_up(L.cand_cap, 16);
        L.out_off = out;
        L.out_cap = L.quota + 64;
        out += L.out_cap;
        L.vec_end = L.w / 32 * 32;
        L.unr_end = L.vec_end + (L.w - L.vec_end) / 4 * 4;
        L.tabx_off = tab;
        tab += L.w;
        L.taby_off = tab;
        tab += L.h;
    }
    C.tiles = tiles, C.cand_total = cand, C.out_total = out;
    W->pyr_slot = align_up((int)pyr, 256) + 256, W->blur_slot = align_up((int)blur, 256);
    const int S = c->max_batch;
    W->slots = S;
    CUDA_TRY(c, dmalloc(&W->d_pyr, W->pyr_slot * S));
    CUDA_TRY(c, dmalloc(&W->d_blur, W->blur_slot * S));
    CUDA_TRY(c, dmalloc(&W->d_score, W->blur_slot * S));
    CUDA_TRY(c, dmalloc(&W->d_cand, (size_t)cand * S * sizeof(unsigned)));
    CUDA_TRY(c, dmalloc(&W->d_lr, (size_t)cand * S * 2 * sizeof(unsigned short)));
    CUDA_TRY(c, dmalloc(&W->d_hkey, (size_t)cand * S * sizeof(HKey)));
    CUDA_TRY(c, dmalloc(&W->d_hcount, (size_t)S * RGBDFE_MAX_LEVELS * sizeof(int)));
    CUDA_TRY(c, dmalloc(&W->d_out, (size_t)out * S * sizeof(CvSel)));
    CUDA_TRY(c, dmalloc(&W->d_out_count, (size_t)S * RGBDFE_MAX_LEVELS * sizeof(int)));
    CUDA_TRY(c, dmalloc(&W->d_final, (size_t)out * S * sizeof(int)));
    CUDA_TRY(c, dmalloc(&W->d_final_count, (size_t)S * sizeof(int)));
    CUDA_TRY(c, dmalloc(&W->d_gscratch, (size_t)out * S * sizeof(HKey)));
    std::vector<int> ofs(tab, 0);
    std::vector<unsigned short> c1(tab, 0);
    std::vector<int2> xpad;
    std::vector<PyrGroup> groups;
    W->pyr_nsrc = 0;
    for (int l = 1; l < C.nlevels; l++) {
        const CvLevel &L = C.lv[l], &P = C.lv[l - 1];
        exact_table(P.w, L.w, &ofs[L.tabx_off], &c1[L.tabx_off]);
        exact_table(P.h, L.h, &ofs[L.taby_off], &c1[L.taby_off]);
        // padded forms: a ring pixel is the level's own pixel at the REFLECT_101 coordinate, columns of the row pitch
        // beyond the padded width repeat the last one
        const int pw = L.w + 2 * CV_RING, ph = L.h + 2 * CV_RING, ng = (ph + PYR_ROWS - 1) / PYR_ROWS;
        auto reflect = [](int p, int len) { p = p < 0 ? -p : p; return p >= len ? 2 * (len - 1) - p : p; };
        W->xpad_off[l] = (int)xpad.size(), W->group_off[l] = (int)groups.size();
        for (int px = 0; px < L.pstride; px++) {
            const int x = reflect(std::min(px, pw - 1) - CV_RING, L.w);
            const unsigned k1 = c1[L.tabx_off + x];
            xpad.push_back(make_int2(ofs[L.tabx_off + x], (int)((256u - k1) | k1 << 16)));
        }
        for (int gi = 0; gi < ng; gi++) {
            PyrGroup G;
            memset(&G, 0, sizeof(G));
            for (int j = 0; j < PYR_ROWS; j++) {
                const int y = reflect(std::min(gi * PYR_ROWS + j, ph - 1) - CV_RING, L.h);
                const int oy = ofs[L.taby_off + y];
                unsigned sel = 0;
                for (int k = 0; k < 2; k++) {
                    const int off = (oy + k) * P.pstride;
                    int i = 0;
                    while (i < G.nsrc && G.off[i] != off)
                        i++;
                    if (i == G.nsrc) {
                        if (G.nsrc == PYR_SRC_MAX) {
                            c->set_error("cvorb: scale factor too large for the pyramid kernel (more than 8 source rows per 4 output rows)");
                            return RGBDFE_E_ARG;
                        }
                        G.off[G.nsrc++] = off;
                    }
                    sel |= (unsigned)i << (4 * k);
                }
                G.sel |= sel << (8 * j);
                const unsigned k1 = c1[L.taby_off + y];
                G.beta[j] = (256u - k1) | k1 << 16;
            }
            W->pyr_nsrc = std::max(W->pyr_nsrc, G.nsrc);
            for (int i = G.nsrc; i < PYR_SRC_MAX; i++)
                G.off[i] = G.off[G.nsrc - 1];
            groups.push_back(G);
        }
    }
    CUDA_TRY(c, dmalloc(&W->d_xpad, std::max<size_t>(xpad.size(), 1) * sizeof(int2)));
    CUDA_TRY(c, dmalloc(&W->d_group, std::max<size_t>(groups.size(), 1) * sizeof(PyrGroup)));
    if (!xpad.empty()) {
        CUDA_TRY(c, cudaMemcpy(W->d_xpad, xpad.data(), xpad.size() * sizeof(int2), cudaMemcpyHostToDevice));
        CUDA_TRY(c, cudaMemcpy(W->d_group, groups.data(), groups.size() * sizeof(PyrGroup), cudaMemcpyHostToDevice));
    }
    {
        EncodeTiledFn enc = encode_tiled_fn();
        if (!enc) {
            c->set_error("cuTensorMapEncodeTiled is not available from this driver (TMA staging needs CUDA >= 12)");
            return RGBDFE_E_CUDA;
        }
        std::vector<CUtensorMap> hm(RGBDFE_MAX_LEVELS);
        for (int l = 0; l < C.nlevels; l++) {
            const CvLevel& L = C.lv[l];
            const cuuint64_t dims[3] = { (cuuint64_t)L.pstride, (cuuint64_t)(L.h + 2 * CV_RING), (cuuint64_t)S };
            const cuuint64_t strides[2] = { (cuuint64_t)L.pstride, (cuuint64_t)W->pyr_slot };
            const cuuint32_t box[3] = { CV_PITCH, CV_BOX_ROWS, 1 }, estr[3] = { 1, 1, 1 };
            // the tensor starts at the padded level's first byte (row -32, column -32)
            void* base = W->d_pyr + L.pyr_off - (size_t)CV_RING * L.pstride - CV_RING;
            const CUresult r = enc(&hm[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) {
                c->set_error("cuTensorMapEncodeTiled failed for a cv::ORB pyramid level");
                return RGBDFE_E_CUDA;
            }
        }
        CUDA_TRY(c, dmalloc(&W->d_maps, hm.size() * sizeof(CUtensorMap)));
        CUDA_TRY(c, cudaMemcpy(W->d_maps, hm.data(), hm.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
    }
    if (!W->d_pattern) {
        CUDA_TRY(c, dmalloc(&W->d_pattern, 1024));
        CUDA_TRY(c, cudaMemcpy(W->d_pattern, kBriefPattern, 1024, cudaMemcpyHostToDevice));
    }
    return RGBDFE_OK;
}

// The whole cv::ORB branch for `count` frames whose gray images are reachable from the device through `gray`.
int launch_cvorb(rgbdfe_ctx* c, int count, int w, int h, const ImgPtrs& gray, int gray_stride, const FrameBatch& frames, bool unproject,
    const ImgPtrs& depth, int depth_stride, const FramePrep& prep)
{
    CvOrbWorkspace* W = c->cvorb;
    if (!W || !W->configured)
        return RGBDFE_E_STATE;
    int rc = cvorb_prepare(c, w, h);
    if (rc != RGBDFE_OK)
        return rc;
    const CvCfg& C = W->cfg;
    {
        ProfScope prof(c, KID_PYRAMID);
        for (int l = 0; l < C.nlevels; l++) {
            const CvLevel& L = C.lv[l];
            const int ph = L.h + 2 * CV_RING;
            if (l == 0) {
                dim3 blk(64, 4), grd((L.pstride / 4 + 63) / 64, (ph + 3) / 4, count);
                k_cv_pyr_base<<<grd, blk, 0, c->stream>>>(C, gray, gray_stride, W->d_pyr, W->pyr_slot);
            } else {
                CvPyr V;
                V.ph = ph, V.pstride = L.pstride;
                V.dst_off = L.pyr_off - (unsigned)(CV_RING * L.pstride + CV_RING), V.src_off = C.lv[l - 1].pyr_off;
                V.xpad_off = W->xpad_off[l], V.group_off = W->group_off[l];
                dim3 blk(64, 4), grd((L.pstride / 4 + 63) / 64, (ph + 4 * PYR_ROWS - 1) / (4 * PYR_ROWS), count);
                if (W->pyr_nsrc <= 6)
                    k_cv_pyr_level<6><<<grd, blk, 0, c->stream>>>(V, W->d_pyr, W->pyr_slot, W->d_xpad, W->d_group);
                else
                    k_cv_pyr_level<8><<<grd, blk, 0, c->stream>>>(V, W->d_pyr, W->pyr_slot, W->d_xpad, W->d_group);
            }
            c->launches++;
        }
    }
    {
        ProfScope prof(c, KID_FAST);
        k_cv_fast<<<dim3(C.tiles, count), 256, 0, c->stream>>>(C, W->d_maps, W->d_score, W->blur_slot);
        c->launches++;
    }
    {
        ProfScope prof(c, KID_BLUR);
        k_cv_blur<<<dim3(C.tiles, count), 256, 0, c->stream>>>(C, W->d_maps, W->d_blur, W->blur_slot);
        c->launches++;
    }
    {
        ProfScope prof(c, KID_QUADTREE); // the selection stage takes the quadtree's place
        int kcap = 4096; // ~ 2x the FAST corners a textured level-0 image yields (w * h / 48), within [4096, 24576]
        while (kcap < w * h / 48 && kcap < 24576)
            kcap *= 2;
        kcap = std::min(kcap, 24576);
        const size_t smem = (size_t)kcap * 8 + (size_t)(h + 2) * sizeof(int);
        CUDA_TRY(c, ensure_dyn_smem(c, k_cv_select, smem));
        k_cv_select<<<dim3(C.nlevels, count), CV_SEL_THREADS, smem, c->stream>>>(C, W->d_pyr, W->pyr_slot, W->d_score, W->blur_slot, W->d_cand,
            W->d_lr, W->d_hkey, W->d_hcount, W->d_out, W->d_out_count, kcap);
        k_cv_gather<<<count, 256, 0, c->stream>>>(C, W->d_out, W->d_out_count, W->d_final, W->d_final_count, W->d_gscratch);
        c->launches += 2;
    }
    {
        ProfScope prof(c, KID_DESCRIBE);
        const int warps = std::min(c->kp_cap, C.out_total);
        k_cv_orient_describe<<<dim3((warps + 7) / 8, count), 256, 0, c->stream>>>(C, W->d_pyr, W->pyr_slot, W->d_blur, W->blur_slot, W->d_out,
            W->d_final, W->d_final_count, W->d_pattern, frames, depth, depth_stride, unproject ? 1 : 0, prep);
        c->launches++;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

// Test taps: level image / blurred level (tight w x h) and the Harris-stage list of slot 0 after the last extraction.
int cvorb_stage_level(rgbdfe_ctx* c, int level, int which, uint8_t* dst, int dst_stride, int* w, int* h)
{
    CvOrbWorkspace* W = c->cvorb;
    if (!W || W->cfg.w == 0 || level < 0 || level >= W->cfg.nlevels)
        return RGBDFE_E_STATE;
    const CvLevel& L = W->cfg.lv[level];
    if (w)
        *w = L.w;
    if (h)
        *h = L.h;
    if (dst) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        if (which == 0)
            CUDA_TRY(c, cudaMemcpy2D(dst, dst_stride, W->d_pyr + L.pyr_off, L.pstride, L.w, L.h, cudaMemcpyDeviceToHost));
        else
            CUDA_TRY(c, cudaMemcpy2D(dst, dst_stride, (which == 1 ? W->d_blur : W->d_score) + L.blur_off, L.bstride, L.w, L.h, cudaMemcpyDeviceToHost));
    }
    return RGBDFE_OK;
}

int cvorb_stage_harris(rgbdfe_ctx* c, int level, rgbdfe_keypoint* out, int cap, int* n)
{
    CvOrbWorkspace* W = c->cvorb;
    if (!W || W->cfg.w == 0 || level < 0 || level >= W->cfg.nlevels || !n)
        return RGBDFE_E_STATE;
    const CvLevel& L = W->cfg.lv[level];
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int cnt = 0;
    CUDA_TRY(c, cudaMemcpy(&cnt, W->d_hcount + level, sizeof(int), cudaMemcpyDeviceToHost));
    *n = cnt;
    const int m = std::min(cnt, cap);
    if (m > 0 && out) {
        std::vector<HKey> tmp(m);
        CUDA_TRY(c, cudaMemcpy(tmp.data(), W->d_hkey + L.cand_off, (size_t)m * sizeof(HKey), cudaMemcpyDeviceToHost));
        for (int i = 0; i < m; i++) {
            out[i].x = (float)(tmp[i].xy & 0xffff);
            out[i].y = (float)(tmp[i].xy >> 16);
            out[i].size = 31 * L.scale;
            out[i].angle = -1.f;
            out[i].response = tmp[i].r;
            out[i].octave = level;
            out[i].class_id = -1;
        }
    }
    return RGBDFE_OK;
}

} // namespace rgbdfe
