// Matcher::projectionMatch (reference Features/Matcher.cpp:35-104), scope row N2.
//
// The reference walks F1's keypoints in order; each projects into F2, looks up the keypoints of F2's 64 x 48 grid
// around the projection (Frame::getFeaturesInArea, Core/Frame.cpp:179-229), keeps the closest descriptor among
// those no earlier keypoint has taken, and takes it when the distance is <= TH_HIGH.
//
// Device form, three launches for any number of pairs:
//   k_proj_candidates<false>  one warp per F1 keypoint: projection (float, no FMA, the order cv::gemm's 3x3 path
//                             uses), then all of F2's keypoints against the cell range + window + valid depth +
//                             Hamming <= TH_HIGH tests; counts the survivors
//   k_proj_offsets            exclusive scan of the counts
//   k_proj_candidates<true>   the same walk, survivors stored as sortable keys  dist << 32 | cell << 16 | index
//                             (the reference meets candidates cell by cell, x outer, y inner, in keypoint order
//                             inside a cell, and keeps the first of equal distances: that is the smallest key)
//   k_proj_resolve            one CTA per pair: the serial "first come, first served" claim as a fixpoint -- every
//                             keypoint takes its smallest key whose train index no EARLIER keypoint currently holds;
//                             after k rounds the first k keypoints have their final answer, a round without change
//                             is the serial result.  Conflicts are rare, so it ends after a few rounds.
#include "common.cuh"

#include <algorithm>
#include <vector>

#define PROJ_TH_HIGH 100 // Matcher::TH_HIGH, Features/Matcher.cpp:9
#define PROJ_WARPS 8
#define GRID_ROWS 48 // FRAME_GRID_ROWS, Core/Frame.h:16
#define GRID_COLS 64

struct ProjPair {
    const uint8_t* q_desc;
    const rgbdfe_point3f* q_xyz;
    const uint8_t* q_outlier; // may be null
    const int* nq;
    const uint8_t* t_desc;
    const rgbdfe_point3f* t_xyz;
    const float2* t_pt;
    const int* t_cell;
    const int* nt;
    rgbdfe_projection P;
};

struct ProjWorkspace {
    int pair_cap = 0;
    ProjPair* d_pairs = nullptr;
    int* d_cnt = nullptr;                // [pairs][kp_cap]
    unsigned* d_off = nullptr;           // [pairs * kp_cap + 1]
    unsigned long long* d_keys = nullptr;
    size_t key_cap = 0;
    rgbdfe_dmatch* d_out = nullptr;      // [pairs][kp_cap]
    int* d_n = nullptr;                  // [pairs]
    int* d_rounds = nullptr;             // [pairs] fixpoint rounds of the last call (diagnostic)
    float2* d_pt = nullptr;              // host-array entry point: F2's undistorted points
    int* d_cell = nullptr;
};

void rgbdfe::free_proj_workspace(rgbdfe_ctx* c)
{
    ProjWorkspace* w = c->proj;
    if (!w)
        return;
    cudaFree(w->d_pairs), cudaFree(w->d_cnt), cudaFree(w->d_off), cudaFree(w->d_keys), cudaFree(w->d_out), cudaFree(w->d_n);
    cudaFree(w->d_rounds), cudaFree(w->d_pt), cudaFree(w->d_cell);
    delete w;
    c->proj = nullptr;
}

// (int) of a float the way x86-64 converts it (cvttss2si): INT_MIN for NaN and out-of-range values.  The reference's
// casts of floor()/ceil() results hit that case for absurd radii only, but the cell range then differs from CUDA's
// saturating conversion.
__device__ __forceinline__ int cvtt_x86(float v)
{
    return (v >= 2147483648.0f || v < -2147483648.0f || v != v) ? (int)0x80000000 : (int)v;
}

// A(3x3, row-major) * b + c with the arithmetic of cv::gemm's small-matrix path for CV_32F: float products summed
// left to right, then the addend (Core/Frame.cpp:324, Features/Matcher.cpp:52).  Compiled with -fmad=false.
__device__ __forceinline__ void mat_vec_add(const float* A, float bx, float by, float bz, const float* c, float& ox, float& oy, float& oz)
{
    ox = ((A[0] * bx + A[1] * by) + A[2] * bz) + c[0];
    oy = ((A[3] * bx + A[4] * by) + A[5] * bz) + c[1];
    oz = ((A[6] * bx + A[7] * by) + A[8] * bz) + c[2];
}

// Frame::posInGrid for the host-array entry point (Core/Frame.cpp:231-241).
__global__ void k_proj_cells(const float2* __restrict__ pt, int n, float minX, float minY, float wInv, float hInv, int* __restrict__ cell)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float2 p = pt[i];
    const int gx = (int)roundf((p.x - minX) * wInv);
    const int gy = (int)roundf((p.y - minY) * hInv);
    cell[i] = (gx < 0 || gx >= GRID_COLS || gy < 0 || gy >= GRID_ROWS) ? -1 : gx * GRID_ROWS + gy;
}

template <bool FILL>
__global__ void __launch_bounds__(PROJ_WARPS * 32) k_proj_candidates(const ProjPair* __restrict__ pairs, int kp_cap, int* __restrict__ cnt,
    const unsigned* __restrict__ off, unsigned long long* __restrict__ keys)
{
    const int pair = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * PROJ_WARPS + (threadIdx.x >> 5);
    if (q >= kp_cap)
        return;
    const ProjPair& pp = pairs[pair];
    const int nq = *pp.nq, nt = *pp.nt;
    int total = 0;
    bool live = q < nq;
    float u = 0.f, v = 0.f;
    int c0 = 0, c1 = -1, r0 = 0, r1 = -1;
    const float r = pp.P.th;
    if (live && pp.q_outlier && pp.q_outlier[q])
        live = false; // Matcher.cpp:46
    if (live) {
        const rgbdfe_point3f p = pp.q_xyz[q];
        live = p.z > 0.0f; // :48 isValidObs
        if (live) {
            float wx, wy, wz, xc, yc, zc;
            mat_vec_add(pp.P.Rwc1, p.x, p.y, p.z, pp.P.Ow1, wx, wy, wz);   // Frame::unprojectWorld
            mat_vec_add(pp.P.Rcw2, wx, wy, wz, pp.P.tcw2, xc, yc, zc);     // :52
            const float invzc = 1.0f / zc;
            u = pp.P.fx * xc * invzc + pp.P.cx;
            v = pp.P.fy * yc * invzc + pp.P.cy;
            // :58-67; a NaN projection passes the reference's comparisons and then finds an empty cell range
            live = !(invzc < 0.0f) && !(u < pp.P.min_x || u > pp.P.max_x) && !(v < pp.P.min_y || v > pp.P.max_y) && u == u && v == v;
        }
    }
    if (live) {
        // Frame::getFeaturesInArea, Core/Frame.cpp:185-199
        c0 = max(0, cvtt_x86(floorf((u - pp.P.min_x - r) * pp.P.grid_w_inv)));
        c1 = min(GRID_COLS - 1, cvtt_x86(ceilf((u - pp.P.min_x + r) * pp.P.grid_w_inv)));
        r0 = max(0, cvtt_x86(floorf((v - pp.P.min_y - r) * pp.P.grid_h_inv)));
        r1 = min(GRID_ROWS - 1, cvtt_x86(ceilf((v - pp.P.min_y + r) * pp.P.grid_h_inv)));
        live = c0 < GRID_COLS && c1 >= 0 && r0 < GRID_ROWS && r1 >= 0;
    }
    if (live) {
        const uint4* qd = reinterpret_cast<const uint4*>(pp.q_desc + (size_t)q * 32);
        const uint4 qa = __ldg(qd), qb = __ldg(qd + 1);
        unsigned long long* dst = FILL ? keys + off[(size_t)pair * kp_cap + q] : nullptr;
        for (int t0 = 0; t0 < nt; t0 += 32) {
            const int t = t0 + lane;
            bool hit = false;
            unsigned long long key = 0;
            if (t < nt) {
                const int cell = __ldg(pp.t_cell + t);
                const int gx = cell / GRID_ROWS, gy = cell - gx * GRID_ROWS;
                if (cell >= 0 && gx >= c0 && gx <= c1 && gy >= r0 && gy <= r1) {
                    const float2 pt = __ldg(pp.t_pt + t);
                    if (fabsf(pt.x - u) < r && fabsf(pt.y - v) < r && pp.t_xyz[t].z > 0.0f) { // :221-224, Matcher.cpp:80
                        const uint4* td = reinterpret_cast<const uint4*>(pp.t_desc + (size_t)t * 32);
                        const uint4 ta = __ldg(td), tb = __ldg(td + 1);
                        const int d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) + __popc(qb.x ^ tb.x)
                            + __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
                        hit = d <= PROJ_TH_HIGH;
                        key = ((unsigned long long)d << 32) | ((unsigned)cell << 16) | (unsigned)t;
                    }
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (FILL && hit)
                dst[total + __popc(m & ((1u << lane) - 1u))] = key;
            total += __popc(m);
        }
    }
    if (!FILL && lane == 0)
        cnt[(size_t)pair * kp_cap + q] = total;
}

// Exclusive scan of m counts into off[0..m] (one CTA).
__global__ void __launch_bounds__(1024) k_proj_offsets(const int* __restrict__ cnt, int m, unsigned* __restrict__ off)
{
    __shared__ unsigned s_warp[32];
    const int per = (m + 1023) / 1024;
    const int b = threadIdx.x * per, e = min(m, b + per);
    unsigned sum = 0;
    for (int i = b; i < e; i++)
        sum += (unsigned)cnt[i];
    unsigned inc = sum;
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31) >= d)
            inc += o;
    }
    if ((threadIdx.x & 31) == 31)
        s_warp[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned w = s_warp[threadIdx.x];
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned o = __shfl_up_sync(0xffffffffu, w, d);
            if (threadIdx.x >= d)
                w += o;
        }
        s_warp[threadIdx.x] = w;
    }
    __syncthreads();
    unsigned run = inc - sum + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u);
    for (int i = b; i < e; i++) {
        off[i] = run;
        run += (unsigned)cnt[i];
    }
    if (threadIdx.x == 1023)
        off[m] = s_warp[31];
}

// The serial claim of Matcher.cpp:45-101 for one pair per CTA.  choice[i] packs dist << 16 | train index, -1: none.
__global__ void __launch_bounds__(1024) k_proj_resolve(const ProjPair* __restrict__ pairs, int kp_cap, const unsigned* __restrict__ off,
    const unsigned long long* __restrict__ keys, rgbdfe_dmatch* __restrict__ out, int* __restrict__ n_out, int* __restrict__ rounds_out)
{
    extern __shared__ int s_mem[];
    int* owner = s_mem;            // [kp_cap] smallest query index currently holding a train keypoint
    int* choice = s_mem + kp_cap;  // [kp_cap]
    int* next = choice + kp_cap;   // [kp_cap]
    __shared__ int s_scan[32];
    const int pair = blockIdx.x;
    const ProjPair& pp = pairs[pair];
    const int nq = min(*pp.nq, kp_cap), nt = min(*pp.nt, kp_cap);
    const unsigned* o = off + (size_t)pair * kp_cap;
    for (int i = threadIdx.x; i < nq; i += blockDim.x)
        choice[i] = -1;
    int rounds = 0;
    for (;;) {
        for (int t = threadIdx.x; t < nt; t += blockDim.x)
            owner[t] = 0x7fffffff;
        __syncthreads();
        for (int i = threadIdx.x; i < nq; i += blockDim.x)
            if (choice[i] >= 0)
                atomicMin(&owner[choice[i] & 0xffff], i);
        __syncthreads();
        int changed = 0;
        for (int i = threadIdx.x; i < nq; i += blockDim.x) {
            unsigned long long best = ~0ull;
            for (unsigned k = o[i], e = o[i + 1]; k < e; k++) {
                const unsigned long long key = keys[k];
                if (owner[(unsigned)key & 0xffffu] >= i) // not held by an earlier keypoint (Matcher.cpp:82)
                    best = key < best ? key : best;
            }
            const int c = best == ~0ull ? -1 : (int)((unsigned)(best >> 32) << 16 | ((unsigned)best & 0xffffu));
            changed |= c != choice[i];
            next[i] = c;
        }
        changed = __syncthreads_or(changed);
        rounds++;
        if (!changed)
            break;
        for (int i = threadIdx.x; i < nq; i += blockDim.x)
            choice[i] = next[i];
        __syncthreads();
    }
    // matches in query order: each thread owns a contiguous run of keypoints
    const int per = (nq + blockDim.x - 1) / blockDim.x;
    const int b = threadIdx.x * per, e = min(nq, b + per);
    int mine = 0;
    for (int i = b; i < e; i++)
        mine += choice[i] >= 0;
    int inc = mine;
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31) >= d)
            inc += v;
    }
    if ((threadIdx.x & 31) == 31)
        s_scan[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        int w = s_scan[threadIdx.x];
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, w, d);
            if (threadIdx.x >= d)
                w += v;
        }
        s_scan[threadIdx.x] = w;
    }
    __syncthreads();
    int pos = inc - mine + ((threadIdx.x >> 5) ? s_scan[(threadIdx.x >> 5) - 1] : 0);
    rgbdfe_dmatch* dst = out + (size_t)pair * kp_cap;
    for (int i = b; i < e; i++) {
        const int c = choice[i];
        if (c >= 0) {
            rgbdfe_dmatch m;
            m.queryIdx = i, m.trainIdx = c & 0xffff, m.imgIdx = -1, m.distance = (float)(c >> 16); // cv::DMatch(i1, bestIdx2, bestDist)
            dst[pos++] = m;
        }
    }
    if (threadIdx.x == 0) {
        n_out[pair] = s_scan[31];
        rounds_out[pair] = rounds;
    }
}

static int ensure_proj(rgbdfe_ctx* c, int npairs)
{
    if (!c->proj)
        c->proj = new ProjWorkspace();
    ProjWorkspace* w = c->proj;
    if (npairs <= w->pair_cap)
        return RGBDFE_OK;
    cudaFree(w->d_pairs), cudaFree(w->d_cnt), cudaFree(w->d_off), cudaFree(w->d_out), cudaFree(w->d_n), cudaFree(w->d_rounds);
    w->d_pairs = nullptr, w->d_cnt = nullptr, w->d_off = nullptr, w->d_out = nullptr, w->d_n = nullptr, w->d_rounds = nullptr;
    w->pair_cap = 0;
    const size_t m = (size_t)npairs * c->kp_cap;
    CUDA_TRY(c, cudaMalloc((void**)&w->d_pairs, sizeof(ProjPair) * npairs));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_cnt, sizeof(int) * m));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_off, sizeof(unsigned) * (m + 1)));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_out, sizeof(rgbdfe_dmatch) * m));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_n, sizeof(int) * npairs));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_rounds, sizeof(int) * npairs));
    if (!w->d_pt) {
        CUDA_TRY(c, cudaMalloc((void**)&w->d_pt, sizeof(float2) * c->kp_cap));
        CUDA_TRY(c, cudaMalloc((void**)&w->d_cell, sizeof(int) * c->kp_cap));
    }
    w->pair_cap = npairs;
    return RGBDFE_OK;
}

// Runs the pass over h_pairs[0 .. npairs) and copies the matches back: out[pair * cap_per_pair ..], n[pair].
static int run_projection(rgbdfe_ctx* c, const std::vector<ProjPair>& h_pairs, rgbdfe_dmatch* out, int cap_per_pair, int* n)
{
    const int npairs = (int)h_pairs.size();
    if (c->kp_cap > 65536) {
        c->set_error("projection match: keypoint capacity above 65536 does not fit the candidate key");
        return RGBDFE_E_CAPACITY;
    }
    int rc = ensure_proj(c, npairs);
    if (rc != RGBDFE_OK)
        return rc;
    ProjWorkspace* w = c->proj;
    cudaStream_t s = c->stream;
    const int cap = c->kp_cap;
    const int m = npairs * cap;
    CUDA_TRY(c, cudaMemcpyAsync(w->d_pairs, h_pairs.data(), sizeof(ProjPair) * npairs, cudaMemcpyHostToDevice, s));
    const dim3 grid((cap + PROJ_WARPS - 1) / PROJ_WARPS, npairs);
    k_proj_candidates<false><<<grid, PROJ_WARPS * 32, 0, s>>>(w->d_pairs, cap, w->d_cnt, nullptr, nullptr);
    k_proj_offsets<<<1, 1024, 0, s>>>(w->d_cnt, m, w->d_off);
    unsigned total = 0;
    CUDA_TRY(c, cudaMemcpyAsync(&total, w->d_off + m, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    if ((size_t)total > w->key_cap) { // candidate storage follows the data: th = 5 needs a few keys per keypoint
        cudaFree(w->d_keys);
        w->d_keys = nullptr, w->key_cap = 0;
        const size_t want = std::max<size_t>((size_t)total + total / 2, (size_t)16 * cap);
        CUDA_TRY(c, cudaMalloc((void**)&w->d_keys, sizeof(unsigned long long) * want));
        w->key_cap = want;
    }
    k_proj_candidates<true><<<grid, PROJ_WARPS * 32, 0, s>>>(w->d_pairs, cap, nullptr, w->d_off, w->d_keys);
    const size_t smem = sizeof(int) * 3 * (size_t)cap;
    if (smem > 48 * 1024)
        CUDA_TRY(c, cudaFuncSetAttribute(k_proj_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_proj_resolve<<<npairs, 1024, smem, s>>>(w->d_pairs, cap, w->d_off, w->d_keys, w->d_out, w->d_n, w->d_rounds);
    c->launches += 4;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(n, w->d_n, sizeof(int) * npairs, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    bool over = false;
    for (int p = 0; p < npairs; p++) {
        const int k = std::min(n[p], cap_per_pair);
        over |= n[p] > cap_per_pair;
        if (k > 0 && out)
            CUDA_TRY(c, cudaMemcpyAsync(out + (size_t)p * cap_per_pair, w->d_out + (size_t)p * cap, sizeof(rgbdfe_dmatch) * k, cudaMemcpyDeviceToHost, s));
    }
    CUDA_TRY(c, cudaStreamSynchronize(s));
    return over ? RGBDFE_E_CAPACITY : RGBDFE_OK;
}

__global__ void k_proj_set2(int* a, int va, int* b, int vb)
{
    *a = va;
    *b = vb;
}

extern "C" {

int rgbdfe_projection_match(rgbdfe_ctx* c, const uint8_t* q_desc, int nq, const rgbdfe_point3f* q_xyz, const uint8_t* q_outlier,
    const uint8_t* t_desc, int nt, const rgbdfe_point3f* t_xyz, const float* t_pts_un, const rgbdfe_projection* proj, rgbdfe_dmatch* out, int cap,
    int* n)
{
    if (!c || !n || !proj || nq < 0 || nt < 0)
        return RGBDFE_E_ARG;
    *n = 0;
    if (nq > c->kp_cap || nt > c->kp_cap) {
        c->set_error("projection match: keypoint count exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    if (nq == 0 || nt == 0)
        return RGBDFE_OK;
    if (!q_desc || !t_desc || !q_xyz || !t_xyz || !t_pts_un)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    int rc = ensure_proj(c, 1);
    if (rc != RGBDFE_OK)
        return rc;
    ProjWorkspace* w = c->proj;
    rgbdfe_frame *Q = c->scratch, *T = c->scratch2;
    cudaStream_t s = c->stream;
    CUDA_TRY(c, cudaMemcpyAsync(Q->desc, q_desc, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(T->desc, t_desc, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(Q->xyz, q_xyz, (size_t)nq * 12, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(T->xyz, t_xyz, (size_t)nt * 12, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(w->d_pt, t_pts_un, (size_t)nt * 8, cudaMemcpyHostToDevice, s));
    if (q_outlier)
        CUDA_TRY(c, cudaMemcpyAsync(Q->outlier, q_outlier, (size_t)nq, cudaMemcpyHostToDevice, s));
    k_proj_set2<<<1, 1, 0, s>>>(Q->n, nq, T->n, nt);
    k_proj_cells<<<(nt + 255) / 256, 256, 0, s>>>(w->d_pt, nt, proj->min_x, proj->min_y, proj->grid_w_inv, proj->grid_h_inv, w->d_cell);
    c->launches += 2;
    std::vector<ProjPair> hp(1);
    hp[0] = ProjPair{ Q->desc, Q->xyz, q_outlier ? Q->outlier : nullptr, Q->n, T->desc, T->xyz, w->d_pt, w->d_cell, T->n, *proj };
    return run_projection(c, hp, out, cap, n);
}

int rgbdfe_frames_projection_match(rgbdfe_ctx* c, rgbdfe_frame* const* f1, rgbdfe_frame* const* f2, const rgbdfe_projection* proj, int npairs,
    rgbdfe_dmatch* out, int cap_per_pair, int* n)
{
    if (!c || !f1 || !f2 || !proj || !n || npairs < 0 || cap_per_pair < 0)
        return RGBDFE_E_ARG;
    if (npairs == 0)
        return RGBDFE_OK;
    cudaSetDevice(c->device);
    std::vector<ProjPair> hp(npairs);
    for (int p = 0; p < npairs; p++) {
        const rgbdfe_frame *a = f1[p], *b = f2[p];
        if (!a || !b)
            return RGBDFE_E_ARG;
        if (a->cap > c->kp_cap || b->cap > c->kp_cap) {
            c->set_error("projection match: frame belongs to a context with a larger keypoint capacity");
            return RGBDFE_E_CAPACITY;
        }
        hp[p] = ProjPair{ a->desc, a->xyz, a->outlier, a->n, b->desc, b->xyz, b->pt_un, b->cell, b->n, proj[p] };
    }
    return run_projection(c, hp, out, cap_per_pair, n);
}

int rgbdfe_projection_rounds(rgbdfe_ctx* c, int* rounds, int npairs)
{
    if (!c || !rounds || !c->proj || npairs > c->proj->pair_cap)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaMemcpy(rounds, c->proj->d_rounds, sizeof(int) * npairs, cudaMemcpyDeviceToHost));
    return RGBDFE_OK;
}

} // extern "C"
