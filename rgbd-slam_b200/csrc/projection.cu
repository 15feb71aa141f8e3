// Matcher::projectionMatch (reference Features/Matcher.cpp:35-104), scope row N2.
//
// The reference walks F1's keypoints in order; each projects into F2, looks up the keypoints of F2's 64 x 48 grid
// around the projection (Frame::getFeaturesInArea, Core/Frame.cpp:179-229), keeps the closest descriptor among
// those no earlier keypoint has taken, and takes it when the distance is <= TH_HIGH.
//
// Device form, five launches for any number of pairs:
//   k_proj_grid               one CTA per pair: F2's mGrid as a CSR table (cell -> keypoint indices with a valid
//                             depth), counting sort in shared memory; the order inside a cell is free, see below
//   k_proj_candidates<false>  one warp per F1 keypoint: projection (float, no FMA, the order cv::gemm's 3x3 path
//                             uses), then the keypoints of the cell range -- one contiguous CSR run per grid
//                             column -- against the window and Hamming <= TH_HIGH tests; counts the survivors
//   k_proj_offsets            exclusive scan of each pair's counts; the host reads the totals and sizes the key store
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
#define GRID_CELLS (GRID_COLS * GRID_ROWS)

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
    int* d_gstart = nullptr;             // [pairs][GRID_CELLS + 1] CSR of F2's grid
    int* d_gitems = nullptr;             // [pairs][kp_cap]
    int* d_cnt = nullptr;                // [pairs][kp_cap]
    unsigned* d_off = nullptr;           // [pairs][kp_cap + 1], the pair's total last
    unsigned long long* d_keys = nullptr; // [pairs][key_stride]
    size_t key_stride = 0, key_cap = 0;
    unsigned* h_tot = nullptr;           // pinned [pairs]
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
    cudaFree(w->d_rounds), cudaFree(w->d_pt), cudaFree(w->d_cell), cudaFree(w->d_gstart), cudaFree(w->d_gitems);
    if (w->h_tot)
        cudaFreeHost(w->h_tot);
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

// Exclusive prefix of v over the 1024 threads of the CTA; *total receives the sum.
__device__ __forceinline__ int block_scan_1024(int v, int* s_warp, int* total)
{
    int inc = v;
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31) >= d)
            inc += o;
    }
    if ((threadIdx.x & 31) == 31)
        s_warp[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        int w = s_warp[threadIdx.x];
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, w, d);
            if (threadIdx.x >= d)
                w += o;
        }
        s_warp[threadIdx.x] = w;
    }
    __syncthreads();
    *total = s_warp[31];
    return inc - v + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0);
}

// Frame::assignFeaturesToGrid (Core/Frame.cpp:75-89) of F2 as a CSR table.  Keypoints without a valid depth are left
// out (Matcher.cpp:80 skips them); the order inside a cell does not matter because the candidate key carries it.
__global__ void __launch_bounds__(1024) k_proj_grid(const ProjPair* __restrict__ pairs, int kp_cap, int* __restrict__ gstart, int* __restrict__ gitems)
{
    __shared__ int s_cnt[GRID_CELLS];
    __shared__ int s_warp[32];
    const ProjPair& pp = pairs[blockIdx.x];
    const int nt = min(*pp.nt, kp_cap);
    int* start = gstart + (size_t)blockIdx.x * (GRID_CELLS + 1);
    int* items = gitems + (size_t)blockIdx.x * kp_cap;
    for (int i = threadIdx.x; i < GRID_CELLS; i += 1024)
        s_cnt[i] = 0;
    __syncthreads();
    for (int t = threadIdx.x; t < nt; t += 1024) {
        const int cell = pp.t_cell[t];
        if (cell >= 0 && pp.t_xyz[t].z > 0.0f)
            atomicAdd(&s_cnt[cell], 1);
    }
    __syncthreads();
    const int c3 = threadIdx.x * 3; // GRID_CELLS == 3 * 1024
    const int a = s_cnt[c3], b = s_cnt[c3 + 1], d = s_cnt[c3 + 2];
    int total;
    const int ex = block_scan_1024(a + b + d, s_warp, &total);
    start[c3] = ex, start[c3 + 1] = ex + a, start[c3 + 2] = ex + a + b;
    s_cnt[c3] = ex, s_cnt[c3 + 1] = ex + a, s_cnt[c3 + 2] = ex + a + b; // now the fill cursors
    if (threadIdx.x == 0)
        start[GRID_CELLS] = total;
    __syncthreads();
    for (int t = threadIdx.x; t < nt; t += 1024) {
        const int cell = pp.t_cell[t];
        if (cell >= 0 && pp.t_xyz[t].z > 0.0f)
            items[atomicAdd(&s_cnt[cell], 1)] = t;
    }
}

template <bool FILL>
__global__ void __launch_bounds__(PROJ_WARPS * 32) k_proj_candidates(const ProjPair* __restrict__ pairs, int kp_cap,
    const int* __restrict__ gstart, const int* __restrict__ gitems, int* __restrict__ cnt, const unsigned* __restrict__ off,
    unsigned long long* __restrict__ keys, size_t key_stride)
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
        unsigned long long* dst = FILL ? keys + (size_t)pair * key_stride + off[(size_t)pair * (kp_cap + 1) + q] : nullptr;
        const int* start = gstart + (size_t)pair * (GRID_CELLS + 1);
        const int* items = gitems + (size_t)pair * kp_cap;
        for (int gx = c0; gx <= c1; gx++) {
            // cells (gx, r0 .. r1) follow each other in the table: one run of keypoints per grid column
            const int kb = __ldg(start + gx * GRID_ROWS + r0), ke = __ldg(start + gx * GRID_ROWS + r1 + 1);
            for (int k0 = kb; k0 < ke; k0 += 32) {
                const int k = k0 + lane;
                bool hit = false;
                unsigned long long key = 0;
                if (k < ke) {
                    const int t = __ldg(items + k);
                    const float2 pt = __ldg(pp.t_pt + t);
                    if (fabsf(pt.x - u) < r && fabsf(pt.y - v) < r) { // Core/Frame.cpp:221-224
                        const uint4* td = reinterpret_cast<const uint4*>(pp.t_desc + (size_t)t * 32);
                        const uint4 ta = __ldg(td), tb = __ldg(td + 1);
                        const int d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) + __popc(qb.x ^ tb.x)
                            + __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
                        hit = d <= PROJ_TH_HIGH;
                        key = ((unsigned long long)d << 32) | ((unsigned)__ldg(pp.t_cell + t) << 16) | (unsigned)t;
                    }
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (FILL && hit)
                    dst[total + __popc(m & ((1u << lane) - 1u))] = key;
                total += __popc(m);
            }
        }
    }
    if (!FILL && lane == 0)
        cnt[(size_t)pair * kp_cap + q] = total;
}

// Exclusive scan of a pair's candidate counts: off[pair][0 .. kp_cap], the pair's total in off[pair][kp_cap].
__global__ void __launch_bounds__(1024) k_proj_offsets(const int* __restrict__ cnt, int kp_cap, unsigned* __restrict__ off)
{
    __shared__ int s_warp[32];
    const int* c = cnt + (size_t)blockIdx.x * kp_cap;
    unsigned* o = off + (size_t)blockIdx.x * (kp_cap + 1);
    const int per = (kp_cap + 1023) / 1024;
    const int b = threadIdx.x * per, e = min(kp_cap, b + per);
    int sum = 0;
    for (int i = b; i < e; i++)
        sum += c[i];
    int total;
    int run = block_scan_1024(sum, s_warp, &total);
    for (int i = b; i < e; i++) {
        o[i] = (unsigned)run;
        run += c[i];
    }
    if (threadIdx.x == 0)
        o[kp_cap] = (unsigned)total;
}

// The serial claim of Matcher.cpp:45-101 for one pair per CTA.  choice[i] packs dist << 16 | train index, -1: none.
__global__ void __launch_bounds__(1024) k_proj_resolve(const ProjPair* __restrict__ pairs, int kp_cap, const unsigned* __restrict__ off,
    const unsigned long long* __restrict__ keys_all, size_t key_stride, rgbdfe_dmatch* __restrict__ out, int* __restrict__ n_out, int* __restrict__ rounds_out)
{
    extern __shared__ int s_mem[];
    int* owner = s_mem;            // [kp_cap] smallest query index currently holding a train keypoint
    int* choice = s_mem + kp_cap;  // [kp_cap]
    int* next = choice + kp_cap;   // [kp_cap]
    __shared__ int s_scan[32];
    const int pair = blockIdx.x;
    const ProjPair& pp = pairs[pair];
    const int nq = min(*pp.nq, kp_cap), nt = min(*pp.nt, kp_cap);
    const unsigned* o = off + (size_t)pair * (kp_cap + 1);
    const unsigned long long* keys = keys_all + (size_t)pair * key_stride;
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
    cudaFree(w->d_gstart), cudaFree(w->d_gitems);
    w->d_pairs = nullptr, w->d_cnt = nullptr, w->d_off = nullptr, w->d_out = nullptr, w->d_n = nullptr, w->d_rounds = nullptr;
    w->d_gstart = nullptr, w->d_gitems = nullptr;
    w->pair_cap = 0;
    const size_t m = (size_t)npairs * c->kp_cap;
    CUDA_TRY(c, cudaMalloc((void**)&w->d_pairs, sizeof(ProjPair) * npairs));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_gstart, sizeof(int) * (size_t)npairs * (GRID_CELLS + 1)));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_gitems, sizeof(int) * m));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_cnt, sizeof(int) * m));
    CUDA_TRY(c, cudaMalloc((void**)&w->d_off, sizeof(unsigned) * (m + npairs)));
    if (w->h_tot)
        cudaFreeHost(w->h_tot);
    w->h_tot = nullptr;
    CUDA_TRY(c, cudaMallocHost((void**)&w->h_tot, sizeof(unsigned) * npairs));
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
    CUDA_TRY(c, cudaMemcpyAsync(w->d_pairs, h_pairs.data(), sizeof(ProjPair) * npairs, cudaMemcpyHostToDevice, s));
    const dim3 grid((cap + PROJ_WARPS - 1) / PROJ_WARPS, npairs);
    k_proj_grid<<<npairs, 1024, 0, s>>>(w->d_pairs, cap, w->d_gstart, w->d_gitems);
    k_proj_candidates<false><<<grid, PROJ_WARPS * 32, 0, s>>>(w->d_pairs, cap, w->d_gstart, w->d_gitems, w->d_cnt, nullptr, nullptr, 0);
    k_proj_offsets<<<npairs, 1024, 0, s>>>(w->d_cnt, cap, w->d_off);
    CUDA_TRY(c, cudaMemcpy2DAsync(w->h_tot, sizeof(unsigned), w->d_off + cap, sizeof(unsigned) * (cap + 1), sizeof(unsigned), npairs,
        cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    size_t most = 0;
    for (int p = 0; p < npairs; p++)
        most = std::max<size_t>(most, w->h_tot[p]);
    if (most > w->key_stride || (size_t)npairs * w->key_stride > w->key_cap) { // candidate storage follows the data: th = 5 needs a key or two per keypoint
        const size_t stride = std::max<size_t>(std::max(most + most / 2, w->key_stride), (size_t)4 * cap);
        if ((size_t)npairs * stride > w->key_cap) {
            cudaFree(w->d_keys);
            w->d_keys = nullptr, w->key_cap = 0;
            CUDA_TRY(c, cudaMalloc((void**)&w->d_keys, sizeof(unsigned long long) * npairs * stride));
            w->key_cap = (size_t)npairs * stride;
        }
        w->key_stride = stride;
    }
    k_proj_candidates<true><<<grid, PROJ_WARPS * 32, 0, s>>>(w->d_pairs, cap, w->d_gstart, w->d_gitems, nullptr, w->d_off, w->d_keys, w->key_stride);
    const size_t smem = sizeof(int) * 3 * (size_t)cap;
    if (smem > 48 * 1024)
        CUDA_TRY(c, cudaFuncSetAttribute(k_proj_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_proj_resolve<<<npairs, 1024, smem, s>>>(w->d_pairs, cap, w->d_off, w->d_keys, w->key_stride, w->d_out, w->d_n, w->d_rounds);
    c->launches += 5;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(n, w->d_n, sizeof(int) * npairs, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    bool over = false;
    int widest = 0;
    for (int p = 0; p < npairs; p++) {
        widest = std::max(widest, std::min(n[p], cap_per_pair));
        over |= n[p] > cap_per_pair;
    }
    if (widest > 0 && out) // one strided copy: the first `widest` matches of every pair
        CUDA_TRY(c, cudaMemcpy2D(out, sizeof(rgbdfe_dmatch) * cap_per_pair, w->d_out, sizeof(rgbdfe_dmatch) * cap, sizeof(rgbdfe_dmatch) * widest, npairs,
            cudaMemcpyDeviceToHost));
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
