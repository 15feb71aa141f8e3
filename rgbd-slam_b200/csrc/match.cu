// K7 k_knn2_hamming + K8 k_ratio_filter — Matcher::match (Features/Matcher.cpp:106-139):
// cv::BFMatcher(NORM_HAMMING)::knnMatch(k=2), ratio test, unique-train claim,
// outlier / valid-depth filters.
//
// K7 is the one dense contraction of the path: hamming(q, t) = |q| + |t| - 2 * (q . t) with the descriptors as
// 256-vectors over {0, 1}, and q . t over all (query, train) pairs is a matrix product.  It runs on the tensor
// pipe through the warp-level u8 MMA m16n8k32 (IMMA.16832 on sm_100a: measured 0.5 per clock and SM, i.e. 8
// descriptor pairs per clock and SM against 2.2 for the POPC/LOP3 formulation on the integer pipes).  Bits are
// expanded to bytes once: a warp keeps 32 queries (2 m-tiles x 8 k-steps, 64 registers) stationary, the CTA
// expands each tile of 128 train descriptors into shared memory in exactly the order the B fragments are read
// (conflict-free 8-byte loads), and the accumulator fragment goes straight into a per-thread top-2 (strict <,
// train index ascending), merged across the 4 lanes of a row at the end.  The k order inside the product is
// free, so lane (g, t) simply owns bits [64t, 64t + 64) of its query rows and of its train column, 8 per k-step.
// (The .b1 .and.popc form of the instruction would take the packed bits directly, but ptxas lowers it to a
// subroutine of 24 MOVM transposes + 8 IMMA per call: slower than the integer pipes.)
// Neighbour order is lexicographic (distance, trainIdx) as in OpenCV.
#include <algorithm>
#include <mutex>

#include "match_filter.cuh"

namespace rgbdfe {

#define KNN_THREADS 256
#define KNN_TILE 128      // train descriptors expanded per step (32 KB of shared memory)
#define KNN_MT 2          // m-tiles (16 queries each) per warp
#define KNN_QPB (KNN_THREADS / 32 * KNN_MT * 16) // queries per CTA = 256
#define KNN_NONE 0x7fffffff

__device__ __forceinline__ void mma_u8(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// bits [4n, 4n + 4) of w as four bytes of 0 / 1
__device__ __forceinline__ unsigned nibble_bytes(unsigned w, int n) { return (((w >> (4 * n)) & 0xfu) * 0x00204081u) & 0x01010101u; }

// A candidate is one sortable word: (|t| - 2 q.t + 256) << 16 | train index, so that unsigned order = lexicographic
// (distance, trainIdx) for a fixed query.  m0 <= m1 are the two smallest seen; pushing is three min/max.
#define KNN_KEY_BIAS 256
#define KNN_EMPTY 0x7fffffffu  // no candidate yet
#define KNN_PAD 0x7ffe0000u    // base word of a column past the end of the train set (never beats a real one)
__device__ __forceinline__ void top2_push(unsigned p, unsigned& m0, unsigned& m1)
{
    m1 = min(m1, max(m0, p));
    m0 = min(m0, p);
}

// partial: [pair][split][q] -> int4 (d0, i0, d1, i1)
__global__ void __launch_bounds__(KNN_THREADS) k_knn2_hamming(const MatchPair* __restrict__ pairs, int slot0, int4* __restrict__ partial, int q_cap,
    int splits)
{
    const int pair = slot0 + blockIdx.z;
    const MatchPair P = pairs[pair];
    const int nq = *P.nq, nt = *P.nt;
    const int q_base = blockIdx.x * KNN_QPB;
    if (q_base >= nq)
        return;
    // expanded train tile: [n-tile of 8 columns][k-step][lane = column * 4 + quarter] -> (b0, b1)
    __shared__ __align__(16) uint2 s_b[KNN_TILE / 8][8][32];
    __shared__ __align__(8) unsigned s_base[KNN_TILE]; // (|t| + 256) << 16 | train index
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3; // fragment row / column and 64-bit quarter of the descriptor
    // stationary operand: bits [64t, 64t + 64) of rows g and g + 8 of each m-tile as bytes, 8 per k-step; |q| per row
    unsigned a[KNN_MT][8][4];
    int qpop[KNN_MT][2];
#pragma unroll
    for (int m = 0; m < KNN_MT; m++) {
        const int r0 = q_base + (warp * KNN_MT + m) * 16 + g, r1 = r0 + 8;
        uint2 w0 = make_uint2(0u, 0u), w1 = w0;
        if (r0 < nq)
            w0 = reinterpret_cast<const uint2*>(P.q_desc + (size_t)r0 * 32)[t];
        if (r1 < nq)
            w1 = reinterpret_cast<const uint2*>(P.q_desc + (size_t)r1 * 32)[t];
#pragma unroll
        for (int s = 0; s < 8; s++) {
            const unsigned v0 = s < 4 ? w0.x : w0.y, v1 = s < 4 ? w1.x : w1.y;
            a[m][s][0] = nibble_bytes(v0, 2 * (s & 3)), a[m][s][2] = nibble_bytes(v0, 2 * (s & 3) + 1);
            a[m][s][1] = nibble_bytes(v1, 2 * (s & 3)), a[m][s][3] = nibble_bytes(v1, 2 * (s & 3) + 1);
        }
        int p0 = __popc(w0.x) + __popc(w0.y), p1 = __popc(w1.x) + __popc(w1.y);
        p0 += __shfl_xor_sync(0xffffffffu, p0, 1), p1 += __shfl_xor_sync(0xffffffffu, p1, 1);
        p0 += __shfl_xor_sync(0xffffffffu, p0, 2), p1 += __shfl_xor_sync(0xffffffffu, p1, 2);
        qpop[m][0] = p0, qpop[m][1] = p1;
    }
    unsigned m0[KNN_MT][2], m1[KNN_MT][2];
#pragma unroll
    for (int m = 0; m < KNN_MT; m++)
#pragma unroll
        for (int h = 0; h < 2; h++)
            m0[m][h] = m1[m][h] = KNN_EMPTY;
    // train range of this split, in whole tiles
    const int tiles = (nt + KNN_TILE - 1) / KNN_TILE;
    const int tps = (tiles + splits - 1) / splits;
    const int t_begin = min(blockIdx.y * tps * KNN_TILE, nt), t_end = min((blockIdx.y + 1) * tps * KNN_TILE, nt);
    for (int t0 = t_begin; t0 < t_end; t0 += KNN_TILE) {
        const int cnt = min(KNN_TILE, t_end - t0);
        __syncthreads();
        // expansion: work item = (column j, quarter tq), 512 per tile; consecutive lanes write consecutive 8-byte slots
        for (int wi = tid; wi < KNN_TILE * 4; wi += KNN_THREADS) {
            const int j = wi >> 2, tq = wi & 3;
            uint2 w = make_uint2(0u, 0u);
            if (j < cnt)
                w = reinterpret_cast<const uint2*>(P.t_desc + (size_t)(t0 + j) * 32)[tq];
#pragma unroll
            for (int s = 0; s < 8; s++) {
                const unsigned v = s < 4 ? w.x : w.y;
                s_b[j >> 3][s][wi & 31] = make_uint2(nibble_bytes(v, 2 * (s & 3)), nibble_bytes(v, 2 * (s & 3) + 1));
            }
            int pop = __popc(w.x) + __popc(w.y);
            pop += __shfl_xor_sync(0xffffffffu, pop, 1);
            pop += __shfl_xor_sync(0xffffffffu, pop, 2);
            if (tq == 0)
                s_base[j] = j < cnt ? ((unsigned)(pop + KNN_KEY_BIAS) << 16 | (unsigned)(t0 + j)) : KNN_PAD;
        }
        __syncthreads();
        for (int n0 = 0; n0 < cnt; n0 += 8) {
            int c[KNN_MT][4];
#pragma unroll
            for (int m = 0; m < KNN_MT; m++)
                c[m][0] = c[m][1] = c[m][2] = c[m][3] = 0;
#pragma unroll
            for (int s = 0; s < 8; s++) {
                const uint2 b = s_b[n0 >> 3][s][lane];
#pragma unroll
                for (int m = 0; m < KNN_MT; m++)
                    mma_u8(c[m], a[m][s][0], a[m][s][1], a[m][s][2], a[m][s][3], b.x, b.y);
            }
            const uint2 base = *reinterpret_cast<const uint2*>(&s_base[n0 + 2 * t]); // this thread's two columns
#pragma unroll
            for (int m = 0; m < KNN_MT; m++) {
                // distance word = base - 2 (q . t) << 16  (+ |q| at the end); padded columns have q . t = 0
                top2_push(base.x - ((unsigned)c[m][0] << 17), m0[m][0], m1[m][0]);
                top2_push(base.y - ((unsigned)c[m][1] << 17), m0[m][0], m1[m][0]);
                top2_push(base.x - ((unsigned)c[m][2] << 17), m0[m][1], m1[m][1]);
                top2_push(base.y - ((unsigned)c[m][3] << 17), m0[m][1], m1[m][1]);
            }
        }
    }
    // the 4 lanes of a row hold disjoint column subsets: two butterfly merges, then lane t == 0 owns the result
#pragma unroll
    for (int m = 0; m < KNN_MT; m++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            unsigned a0 = m0[m][h], a1 = m1[m][h];
#pragma unroll
            for (int x = 1; x <= 2; x <<= 1) {
                const unsigned o0 = __shfl_xor_sync(0xffffffffu, a0, x), o1 = __shfl_xor_sync(0xffffffffu, a1, x);
                const unsigned lo = min(a0, o0);
                a1 = min(max(a0, o0), min(a1, o1));
                a0 = lo;
            }
            const int q = q_base + (warp * KNN_MT + m) * 16 + g + 8 * h;
            if (t == 0 && q < nq) {
                int4 r = make_int4(KNN_NONE, -1, KNN_NONE, -1);
                if (a0 < KNN_PAD)
                    r.x = qpop[m][h] + (int)(a0 >> 16) - KNN_KEY_BIAS, r.y = (int)(a0 & 0xffffu);
                if (a1 < KNN_PAD)
                    r.z = qpop[m][h] + (int)(a1 >> 16) - KNN_KEY_BIAS, r.w = (int)(a1 & 0xffffu);
                partial[((size_t)pair * splits + blockIdx.y) * q_cap + q] = r;
            }
        }
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

// Train-range splits per (pair, query block): enough CTAs for about six waves of the two CTAs an SM holds, so that
// the last wave is full (a 64-pair chain alone gives 512 equal CTAs on 296 slots: 1.7 waves), within the partial
// buffer (max(npairs, 64 * knn_splits) rows).
int knn_splits_for(const rgbdfe_ctx* c, int npairs)
{
    const int qpb = c->knn_umma ? 128 : KNN_QPB; // queries per CTA of the selected kernel
    const int qblocks = (c->kp_cap + qpb - 1) / qpb;
    // the persistent kernel (knn_umma == 2) walks work items with a grid stride: splitting only has to give every SM a couple of items
    const int want = c->knn_umma == 2 ? (2 * 148 + qblocks * npairs - 1) / (qblocks * npairs)
                                      : (c->knn_waves * 2 * 148 + qblocks * npairs - 1) / (qblocks * npairs);
    const int room = std::max(npairs, 64 * c->knn_splits) / npairs;
    return std::max(1, std::min(std::min(want, c->knn_splits), room));
}

// K7 for pair slots [slot0, slot0 + npairs)
int launch_knn(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits)
{
    if (npairs <= 0)
        return RGBDFE_OK;
    if (c->kp_cap >= 65536) {
        c->set_error("kNN: train index does not fit the 16-bit field of the candidate word");
        return RGBDFE_E_CAPACITY;
    }
    ProfScope prof(c, KID_KNN);
    c->launches++;
    if (c->knn_umma) // the tcgen05 form (match_umma.cu); the mma.sync form below stays for A/B measurements
        return launch_knn_umma(c, d_pairs, slot0, npairs, max_nq, splits);
    dim3 grd((max_nq + KNN_QPB - 1) / KNN_QPB, splits, npairs);
    k_knn2_hamming<<<grd, KNN_THREADS, 0, c->stream>>>(d_pairs, slot0, c->d_partial, c->kp_cap, splits);
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

// ---- rgbdfe_match_knn2_noalloc: Matcher::match without any allocation ----------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  The scratch of K7 / K8 —
// the pair descriptor, the two device counts and the [split][query] top-2 partials — lives in ONE file-scope
// device array that every GPU loading the library holds until exit.  Its name is a host-side stand-in in host
// code: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
// All calls share the array: calls that overlap on the GPU (different streams without an ordering) race.
#define NOALLOC_MAX_DESC 4096 // most query / train descriptors per call
#define NOALLOC_SPLITS 4      // train-range splits of K7 (fixed: the partials are sized for it)
struct NoallocScratch {
    MatchPair pair;
    int nq, nt;
    int pad[46];
    int4 partial[NOALLOC_SPLITS * NOALLOC_MAX_DESC]; // 256 KB
};
static_assert(sizeof(MatchPair) == 64 && sizeof(NoallocScratch) == 256 + 16 * NOALLOC_SPLITS * NOALLOC_MAX_DESC, "scratch layout");
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoallocScratch)];

// Fills the header (everything the two kernels read before writing).  The partials need no clearing: K7 writes
// every [split][q < nq] entry, also for an empty train range, before K8 reads exactly those.
__global__ void k_noalloc_setup(NoallocScratch* s, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt, const uint8_t* q_outlier,
    const rgbdfe_point3f* q_xyz, const rgbdfe_point3f* t_xyz, int* n_out)
{
    s->nq = nq, s->nt = nt;
    s->pair = MatchPair { q_desc, t_desc, &s->nq, &s->nt, q_outlier, q_xyz, t_xyz, nullptr };
    *n_out = 0;
}

static std::mutex g_noalloc_mu;
static NoallocScratch* g_noalloc_ptr[64]; // device address of g_noalloc_scratch, per device ordinal

} // namespace rgbdfe

using namespace rgbdfe;

extern "C" int rgbdfe_match_knn2_noalloc(rgbdfe_ctx* c, void* stream, const uint8_t* q_desc_dev, int nq, const uint8_t* t_desc_dev, int nt,
    const uint8_t* q_outlier_dev, const rgbdfe_point3f* q_xyz_dev, const rgbdfe_point3f* t_xyz_dev, float ratio, int discard_outliers,
    rgbdfe_dmatch* out_dev, int cap, int* n_dev)
{
    if (!c)
        return RGBDFE_E_ARG;
    if (nq < 0 || nt < 0 || cap < 0 || !n_dev || (nq > 0 && (!q_desc_dev || !q_xyz_dev)) || (nt > 0 && (!t_desc_dev || !t_xyz_dev))
        || (cap > 0 && !out_dev)) {
        c->set_error("match_knn2_noalloc: null pointer or negative size");
        return RGBDFE_E_ARG;
    }
    if (nq > NOALLOC_MAX_DESC || nt > NOALLOC_MAX_DESC) {
        c->set_error("match_knn2_noalloc: " + std::to_string(std::max(nq, nt)) + " descriptors exceed the static scratch (at most "
            + std::to_string(NOALLOC_MAX_DESC) + " per side); use rgbdfe_match_knn2");
        return RGBDFE_E_CAPACITY;
    }
    if (c->device < 0 || c->device >= 64) {
        c->set_error("match_knn2_noalloc: device ordinal out of range");
        return RGBDFE_E_ARG;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    NoallocScratch* scr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[c->device]) { // set-up, once per device, with that device current
            void* p = nullptr;
            CUDA_TRY(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[c->device] = static_cast<NoallocScratch*>(p);
        }
        scr = g_noalloc_ptr[c->device];
    }
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c->stream;
    k_noalloc_setup<<<1, 1, 0, s>>>(scr, q_desc_dev, nq, t_desc_dev, nt, q_outlier_dev, q_xyz_dev, t_xyz_dev, n_dev);
    c->launches++;
    if (nq > 0 && nt > 0) {
        const int match_cap = std::min(cap, NOALLOC_MAX_DESC);
        dim3 grd((nq + KNN_QPB - 1) / KNN_QPB, NOALLOC_SPLITS, 1);
        k_knn2_hamming<<<grd, KNN_THREADS, 0, s>>>(&scr->pair, 0, scr->partial, NOALLOC_MAX_DESC, NOALLOC_SPLITS);
        k_ratio_filter<<<1, 256, (size_t)nt * sizeof(int), s>>>(&scr->pair, 0, scr->partial, NOALLOC_MAX_DESC, NOALLOC_SPLITS, ratio, discard_outliers,
            out_dev, match_cap, n_dev, nullptr);
        c->launches += 2;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}
