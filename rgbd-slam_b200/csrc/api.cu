// C ABI of librgbdfe.so (include/rgbdfe.h): context, configuration, host-buffer entry
// points that mirror the reference classes, device-resident frame pipeline, batched
// verification.  No CPU fallback: every entry point runs CUDA kernels on the context's
// stream and reports CUDA errors as RGBDFE_E_CUDA.
#include "common.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <mutex>
#include <new>

using namespace rgbdfe;

namespace {

int cv_round_f(float v) { return (int)lrintf(v); }
int cv_floor_f(float v)
{
    int i = (int)v;
    return i - (i > v);
}

template <class T>
cudaError_t dmalloc(T** p, size_t n)
{
    return cudaMalloc((void**)p, n * sizeof(T) > 0 ? n * sizeof(T) : 1);
}

// cv::resize(INTER_LINEAR) coefficient tables for one level (OpenCV resize.cpp: 11-bit
// fixed point, coordinates (dx+0.5)*scale-0.5 computed in double, narrowed to float).
void resize_tables(int sw, int sh, int dw, int dh, int* xofs, short* xa, int* yofs, short* yb)
{
    const double scale_x = 1. / ((double)dw / sw), scale_y = 1. / ((double)dh / sh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor_f(fx);
        fx -= sx;
        if (sx < 0) {
            fx = 0;
            sx = 0;
        }
        if (sx >= sw - 1) {
            fx = 0;
            sx = sw - 1;
        }
        xofs[dx] = sx;
        xa[dx * 2] = (short)cv_round_f((1.f - fx) * 2048);
        xa[dx * 2 + 1] = (short)cv_round_f(fx * 2048);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor_f(fy);
        fy -= sy;
        yofs[dy] = sy;
        yb[dy * 2] = (short)cv_round_f((1.f - fy) * 2048);
        yb[dy * 2 + 1] = (short)cv_round_f(fy * 2048);
    }
}

// Build the per-level geometry for a w x h image (ORBextractor.cpp:773-782, 613-634) and
// upload it with the resize tables.
int build_level_config(rgbdfe_ctx* c, int w, int h)
{
    if (c->cfg.w == w && c->cfg.h == h && c->cfg.nlevels == c->nlevels && c->cfg.nfeatures == c->nfeatures)
        return RGBDFE_OK;
    if (w > c->max_w || h > c->max_h || w < 64 || h < 64) {
        c->set_error("image size outside [64, max_w] x [64, max_h]");
        return RGBDFE_E_ARG;
    }
    OrbConfig& g = c->cfg;
    memset(&g, 0, sizeof(g));
    g.nlevels = c->nlevels;
    g.nfeatures = c->nfeatures;
    g.iniTh = c->iniTh;
    g.minTh = c->minTh;
    g.w = w;
    g.h = h;
    static const int umax[16] = { 15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3 }; // ORBextractor.cpp:391-405
    memcpy(g.umax, umax, sizeof(umax));
    unsigned pyr = 0, blur = 0;
    int cells = 0, cand = 0, sel = 0, tx = 0, ty = 0, tiles = 0;
    std::vector<int> xofs, yofs;
    std::vector<short> xa, yb;
    std::vector<int2> xpad;
    std::vector<PyrGroup> groups;
    int nsrc_max = 0;
    for (int l = 0; l < g.nlevels; l++) {
        LevelDesc& L = g.lv[l];
        const float s = c->inv_scale[l];
        L.w = cv_round_f((float)w * s);
        L.h = cv_round_f((float)h * s);
        if (L.w < 40 || L.h < 40) {
            c->set_error("pyramid level smaller than 40 px; reduce nlevels");
            return RGBDFE_E_ARG;
        }
        L.pstride = align_up(L.w + 2 * EDGE_THRESHOLD, 64);
        L.bstride = align_up(L.w, 64);
        L.pyr_off = pyr;
        pyr += (unsigned)L.pstride * (L.h + 2 * EDGE_THRESHOLD);
        L.blur_off = blur;
        blur += (unsigned)L.bstride * L.h;
        L.maxBorderX = L.w - EDGE_THRESHOLD + 3;
        L.maxBorderY = L.h - EDGE_THRESHOLD + 3;
        const float width = (float)(L.maxBorderX - 16), height = (float)(L.maxBorderY - 16);
        L.nCols = (int)(width / 30.f);
        L.nRows = (int)(height / 30.f);
        if (L.nCols < 1 || L.nRows < 1) {
            c->set_error("pyramid level too small for a FAST cell");
            return RGBDFE_E_ARG;
        }
        L.wCell = (int)ceilf(width / L.nCols);
        L.hCell = (int)ceilf(height / L.nRows);
        if (L.wCell + 6 > CELL_MAX_DIM || L.hCell + 6 > CELL_MAX_DIM) {
            c->set_error("FAST cell exceeds CELL_MAX_DIM");
            return RGBDFE_E_CAPACITY;
        }
        L.cell_off = cells;
        cells += L.nCols * L.nRows;
        L.cand_off = cand;
        L.cand_cap = L.nCols * L.nRows * CELL_CAP;
        cand += L.cand_cap;
        L.quota = c->quota[l];
        L.sel_off = sel;
        L.sel_cap = L.quota + 16;
        sel += L.sel_cap;
        L.scale = c->scale[l];
        L.patch_size = (int)(31 * c->scale[l]);
        g.max_quota = std::max(g.max_quota, L.quota);
        L.tab_x_off = tx;
        L.tab_y_off = ty;
        g.blur_tile_start[l] = tiles;
        tiles += ((L.w + 63) / 64) * ((L.h + 31) / 32);
        if (l > 0) {
            const LevelDesc& P = g.lv[l - 1];
            xofs.resize(L.w), xa.resize(2 * L.w), yofs.resize(L.h), yb.resize(2 * L.h);
            resize_tables(P.w, P.h, L.w, L.h, xofs.data(), xa.data(), yofs.data(), yb.data());
            // padded forms: a ring pixel is the level's own pixel at the REFLECT_101 coordinate (what copyMakeBorder
            // copies), columns past the padded width (row pitch) repeat the last one
            const int pw = L.w + 2 * EDGE_THRESHOLD, ph = L.h + 2 * EDGE_THRESHOLD;
            const int ng = (ph + PYR_ROWS - 1) / PYR_ROWS;
            xpad.resize(tx + L.pstride);
            groups.resize(ty + ng);
            for (int px = 0; px < L.pstride; px++) {
                const int x = reflect101(std::min(px, pw - 1) - EDGE_THRESHOLD, L.w);
                xpad[tx + px] = make_int2(EDGE_THRESHOLD + xofs[x], (int)((unsigned short)xa[2 * x] | (unsigned)(unsigned short)xa[2 * x + 1] << 16));
            }
            for (int gi = 0; gi < ng; gi++) {
                PyrGroup G;
                memset(&G, 0, sizeof(G));
                for (int j = 0; j < PYR_ROWS; j++) {
                    const int py = std::min(gi * PYR_ROWS + j, ph - 1);
                    const int y = reflect101(py - EDGE_THRESHOLD, L.h);
                    const int r[2] = { std::min(std::max(yofs[y], 0), P.h - 1), std::min(std::max(yofs[y] + 1, 0), P.h - 1) };
                    unsigned sel = 0;
                    for (int k = 0; k < 2; k++) {
                        const int off = (EDGE_THRESHOLD + r[k]) * P.pstride;
                        int i = 0;
                        while (i < G.nsrc && G.off[i] != off)
                            i++;
                        if (i == G.nsrc) {
                            if (G.nsrc == PYR_SRC_MAX) {
                                c->set_error("scale factor too large for the pyramid kernel (more than 8 source rows per 4 output rows)");
                                return RGBDFE_E_ARG;
                            }
                            G.off[G.nsrc++] = off;
                        }
                        sel |= (unsigned)i << (4 * k);
                    }
                    G.sel |= sel << (8 * j);
                    G.beta[j] = (unsigned)(unsigned short)yb[2 * y] | (unsigned)(unsigned short)yb[2 * y + 1] << 16;
                }
                nsrc_max = std::max(nsrc_max, G.nsrc);
                for (int i = G.nsrc; i < PYR_SRC_MAX; i++)
                    G.off[i] = G.off[G.nsrc - 1];
                groups[ty + gi] = G;
            }
            tx += L.pstride;
            ty += ng;
        }
    }
    g.blur_tile_start[g.nlevels] = tiles;
    g.total_cells = cells;
    g.total_cand = cand;
    g.total_sel = sel;
    g.pyr_bytes = pyr;
    g.blur_bytes = blur;
    if ((size_t)pyr > c->pyr_slot || (size_t)blur > c->blur_slot || cells > c->cell_slot || cand > c->cand_slot || sel > c->sel_slot
        || tx > c->tab_x_cap || ty > c->tab_y_cap) {
        c->set_error("internal: workspace too small for this configuration");
        return RGBDFE_E_CAPACITY;
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(c->d_cfg, &g, sizeof(g), cudaMemcpyHostToDevice));
    if (tx > 0) {
        CUDA_TRY(c, cudaMemcpy(c->d_tab_xpad, xpad.data(), tx * sizeof(int2), cudaMemcpyHostToDevice));
        CUDA_TRY(c, cudaMemcpy(c->d_tab_group, groups.data(), ty * sizeof(PyrGroup), cudaMemcpyHostToDevice));
        c->pyr_nsrc = nsrc_max;
    }
    return RGBDFE_OK;
}

int alloc_frame(rgbdfe_ctx* c, rgbdfe_frame** out)
{
    rgbdfe_frame* f = new (std::nothrow) rgbdfe_frame();
    if (!f)
        return RGBDFE_E_ARG;
    memset(f, 0, sizeof(*f));
    f->cap = c->kp_cap;
    CUDA_TRY(c, dmalloc(&f->kps, f->cap));
    CUDA_TRY(c, dmalloc(&f->desc, (size_t)f->cap * 32));
    CUDA_TRY(c, dmalloc(&f->xyz, f->cap));
    CUDA_TRY(c, dmalloc(&f->outlier, f->cap));
    CUDA_TRY(c, dmalloc(&f->n, 1));
    CUDA_TRY(c, dmalloc(&f->pt_un, f->cap));
    CUDA_TRY(c, dmalloc(&f->cell, f->cap));
    CUDA_TRY(c, cudaMemset(f->n, 0, sizeof(int)));
    CUDA_TRY(c, cudaMemset(f->outlier, 0, f->cap));
    CUDA_TRY(c, cudaMemset(f->xyz, 0, sizeof(rgbdfe_point3f) * f->cap));
    *out = f;
    return RGBDFE_OK;
}

void free_frame(rgbdfe_frame* f)
{
    if (!f)
        return;
    cudaFree(f->kps);
    cudaFree(f->desc);
    cudaFree(f->xyz);
    cudaFree(f->outlier);
    cudaFree(f->n);
    cudaFree(f->pt_un);
    cudaFree(f->cell);
    delete f;
}

} // namespace

namespace {

__global__ void k_set_pair(MatchPair* dst, MatchPair v) { *dst = v; }
// consecutive frames of a tracked sequence -> one MatchPair per (previous, current)
__global__ void k_build_chain_pairs(FrameChain fc, int np, MatchPair* dst)
{
    const int p = threadIdx.x;
    if (p >= np)
        return;
    const rgbdfe_frame &ref = fc.f[p], &cur = fc.f[p + 1];
    dst[p] = MatchPair { ref.desc, cur.desc, ref.n, cur.n, ref.outlier, ref.xyz, cur.xyz, cur.outlier };
}
__global__ void k_set_int(int* dst, int v) { *dst = v; }
// rows [0, min(n, cap)) of every frame's keypoints / descriptors / points into [frame][cap] blocks, word by word
__global__ void k_pack_frames(FrameBatch fb, unsigned* __restrict__ kps, unsigned* __restrict__ desc, unsigned* __restrict__ xyz, int* __restrict__ n_out,
    int cap)
{
    const rgbdfe_frame& f = fb.f[blockIdx.y];
    const int cnt = *f.n;
    const int n = min(max(cnt, 0), min(cap, f.cap));
    if (blockIdx.x == 0 && threadIdx.x == 0)
        n_out[blockIdx.y] = cnt;
    const int t = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    if (kps) {
        const unsigned* src = reinterpret_cast<const unsigned*>(f.kps);
        unsigned* dst = kps + (size_t)blockIdx.y * cap * 7;
        for (int i = t; i < n * 7; i += nt)
            dst[i] = src[i];
    }
    if (desc) {
        const uint4* src = reinterpret_cast<const uint4*>(f.desc);
        uint4* dst = reinterpret_cast<uint4*>(desc + (size_t)blockIdx.y * cap * 8);
        for (int i = t; i < n * 2; i += nt)
            dst[i] = src[i];
    }
    if (xyz) {
        const unsigned* src = reinterpret_cast<const unsigned*>(f.xyz);
        unsigned* dst = xyz + (size_t)blockIdx.y * cap * 3;
        for (int i = t; i < n * 3; i += nt)
            dst[i] = src[i];
    }
}

void free_orb_workspaces(rgbdfe_ctx* c)
{
    cudaFree(c->d_img), cudaFree(c->d_depth), cudaFree(c->d_pyr), cudaFree(c->d_blur);
    cudaFree(c->d_cell_count), cudaFree(c->d_cell_list), cudaFree(c->d_cand), cudaFree(c->d_label);
    cudaFree(c->d_cand_count), cudaFree(c->d_sel), cudaFree(c->d_sel_count);
    cudaFree(c->d_tab_xpad), cudaFree(c->d_tab_group);
    c->d_img = nullptr, c->d_depth = nullptr, c->d_pyr = nullptr, c->d_blur = nullptr;
    c->d_cell_count = nullptr, c->d_cell_list = nullptr, c->d_cand = nullptr, c->d_label = nullptr;
    c->d_cand_count = nullptr, c->d_sel = nullptr, c->d_sel_count = nullptr;
    c->d_tab_xpad = nullptr, c->d_tab_group = nullptr;
}

void free_pair_workspaces(rgbdfe_ctx* c)
{
    for (auto& g : c->chain_graph_cache) // the captured launches name these buffers
        if (g.exec)
            cudaGraphExecDestroy(g.exec);
    c->chain_graph_cache.clear();
    cudaFree(c->d_pairs), cudaFree(c->d_partial), cudaFree(c->d_matches), cudaFree(c->d_sorted), cudaFree(c->d_P1), cudaFree(c->d_P2), cudaFree(c->d_w), cudaFree(c->d_nmatch), cudaFree(c->d_samples);
    cudaFree(c->d_rng_used), cudaFree(c->d_hyp), cudaFree(c->d_hyp_mask), cudaFree(c->d_replay), cudaFree(c->d_inl_pos), cudaFree(c->d_pair_rng), cudaFree(c->d_hyp_ready), cudaFree(c->d_pair_lock);
    c->d_sorted = nullptr, c->d_P1 = nullptr, c->d_P2 = nullptr, c->d_w = nullptr;
    c->d_pairs = nullptr, c->d_partial = nullptr, c->d_matches = nullptr, c->d_nmatch = nullptr, c->d_samples = nullptr;
    c->d_rng_used = nullptr, c->d_hyp = nullptr, c->d_hyp_mask = nullptr, c->d_replay = nullptr, c->d_inl_pos = nullptr, c->d_pair_rng = nullptr, c->d_hyp_ready = nullptr, c->d_pair_lock = nullptr;
    c->pair_cap = 0;
}

int ensure_pairs(rgbdfe_ctx* c, int npairs)
{
    if (npairs <= c->pair_cap)
        return RGBDFE_OK;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    free_pair_workspaces(c);
    const size_t P = (size_t)npairs, K = (size_t)c->kp_cap;
    c->mask_words = (c->kp_cap + 31) / 32;
    const size_t partial_rows = std::max(P, (size_t)64 * c->knn_splits);
    CUDA_TRY(c, dmalloc(&c->d_pairs, P));
    CUDA_TRY(c, dmalloc(&c->d_partial, partial_rows * K));
    CUDA_TRY(c, dmalloc(&c->d_matches, P * K));
    CUDA_TRY(c, dmalloc(&c->d_sorted, P * K));
    CUDA_TRY(c, dmalloc(&c->d_P1, P * K));
    CUDA_TRY(c, dmalloc(&c->d_P2, P * K));
    CUDA_TRY(c, dmalloc(&c->d_w, P * K));
    CUDA_TRY(c, dmalloc(&c->d_nmatch, P));
    CUDA_TRY(c, dmalloc(&c->d_samples, P * MAX_ITERS * 8));
    CUDA_TRY(c, dmalloc(&c->d_rng_used, P * MAX_ITERS));
    CUDA_TRY(c, dmalloc(&c->d_hyp, P * MAX_ITERS));
    CUDA_TRY(c, dmalloc(&c->d_hyp_mask, P * MAX_ITERS * c->mask_words));
    CUDA_TRY(c, dmalloc(&c->d_replay, P));
    CUDA_TRY(c, dmalloc(&c->d_inl_pos, P * K));
    CUDA_TRY(c, dmalloc(&c->d_pair_rng, P));
    CUDA_TRY(c, dmalloc(&c->d_hyp_ready, P * MAX_ITERS));
    CUDA_TRY(c, dmalloc(&c->d_pair_lock, P));
    c->pair_cap = npairs;
    return RGBDFE_OK;
}

} // namespace

extern "C" {

int rgbdfe_create(rgbdfe_ctx** out, int device, int max_w, int max_h, int max_feats, int max_batch)
{
    if (!out || max_w < 64 || max_h < 64 || max_w > 4095 || max_h > 4095 || max_feats < 1 || max_batch < 1 || max_batch > 64)
        return RGBDFE_E_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return RGBDFE_E_CUDA; // no CUDA device: there is no CPU fallback
    rgbdfe_ctx* c = new (std::nothrow) rgbdfe_ctx();
    if (!c)
        return RGBDFE_E_ARG;
    c->device = device;
    c->max_w = max_w, c->max_h = max_h, c->max_feats = max_feats, c->max_batch = max_batch;
    c->kp_cap = align_up(max_feats + 64, 32);
    memset(&c->cfg, 0, sizeof(c->cfg));
    auto fail = [&](const char* what) {
        fprintf(stderr, "rgbdfe_create: %s: %s\n", what, cudaGetErrorString(cudaGetLastError()));
        delete c;
        return RGBDFE_E_CUDA;
    };
    if (cudaSetDevice(device) != cudaSuccess)
        return fail("cudaSetDevice");
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess)
        return fail("cudaStreamCreate");
    if (dmalloc(&c->d_cfg, 1) != cudaSuccess || dmalloc(&c->d_rng, 1) != cudaSuccess || dmalloc(&c->d_latch, 1) != cudaSuccess
        || dmalloc(&c->d_const_n, 16) != cudaSuccess)
        return fail("cudaMalloc");
    c->result_cap = 4096;
    if (dmalloc(&c->d_results, c->result_cap) != cudaSuccess || cudaMallocHost((void**)&c->h_results, sizeof(rgbdfe_se3_result) * c->result_cap) != cudaSuccess)
        return fail("cudaMalloc results");
    cudaMemset(c->d_latch, 0, sizeof(DepthLatch));
    // raster covariances of errorFunction2 (SolverSE3.cpp:218-225)
    const double cam_angle_x = 58.0 / 180.0 * M_PI, cam_angle_y = 45.0 / 180.0 * M_PI;
    const double sx = 3 * tan(cam_angle_x / 640), sy = 3 * tan(cam_angle_y / 480);
    c->rcx = sx * sx;
    c->rcy = sy * sy;
    if (alloc_frame(c, &c->scratch) != RGBDFE_OK || alloc_frame(c, &c->scratch2) != RGBDFE_OK)
        return fail("frame alloc");
    *out = c;
    rgbdfe_rng_seed(c, 1); // srand() default state; the reference seeds with time(NULL) (System/Random.cpp:7-13)
    int rc = rgbdfe_orb_configure(c, std::min(1000, max_feats), 1.2f, 8, 20, 7); // Extractor ctor defaults (Extractor.cpp:21)
    if (rc != RGBDFE_OK) {
        fprintf(stderr, "rgbdfe_create: %s\n", c->err.c_str());
        rgbdfe_destroy(c);
        *out = nullptr;
        return rc;
    }
    return RGBDFE_OK;
}

void rgbdfe_destroy(rgbdfe_ctx* c)
{
    if (!c)
        return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_orb_workspaces(c);
    free_pair_workspaces(c);
    free_proj_workspace(c);
    free_verify_state(c);
    free_cvorb_workspace(c);
    free_fast_state(c);
    free_frame(c->scratch);
    free_frame(c->scratch2);
    if (c->order_evt)
        cudaEventDestroy(c->order_evt);
    if (c->side_stream) {
        cudaStreamDestroy(c->side_stream);
        cudaEventDestroy(c->side_fork);
        cudaEventDestroy(c->side_join);
    }
    if (c->bulk_evt)
        cudaEventDestroy(c->bulk_evt);
    if (c->bulk_stream)
        cudaStreamDestroy(c->bulk_stream);
    cudaFree(c->d_packed), cudaFreeHost(c->h_packed), cudaFree(c->d_dl);
    cudaFree(c->d_dbg), cudaFree(c->d_cfg), cudaFree(c->d_rng), cudaFree(c->d_latch), cudaFree(c->d_const_n), cudaFree(c->d_results);
    cudaFreeHost(c->h_results);
    release_stream(c);
    delete c;
}

const char* rgbdfe_last_error(const rgbdfe_ctx* c) { return c ? c->err.c_str() : "null context"; }
void* rgbdfe_stream(rgbdfe_ctx* c) { return c ? (void*)c->stream : nullptr; }
long long rgbdfe_launch_count(const rgbdfe_ctx* c) { return c ? c->launches : 0; }
int rgbdfe_stream_wait(rgbdfe_ctx* waiter, rgbdfe_ctx* signaler)
{
    if (!waiter || !signaler || waiter->device != signaler->device)
        return RGBDFE_E_ARG;
    if (waiter == signaler)
        return RGBDFE_OK;
    cudaSetDevice(signaler->device);
    if (!signaler->order_evt)
        CUDA_TRY(signaler, cudaEventCreateWithFlags(&signaler->order_evt, cudaEventDisableTiming));
    CUDA_TRY(signaler, cudaEventRecord(signaler->order_evt, signaler->stream));
    CUDA_TRY(waiter, cudaStreamWaitEvent(waiter->stream, signaler->order_evt, 0));
    return RGBDFE_OK;
}

int rgbdfe_set_stream_priority(rgbdfe_ctx* c, int high)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int lo = 0, hi = 0;
    CUDA_TRY(c, cudaDeviceGetStreamPriorityRange(&lo, &hi)); // numerically lower = higher priority
    cudaStream_t s = nullptr;
    CUDA_TRY(c, cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, high ? hi : lo));
    release_stream(c);
    c->stream = s;
    c->owns_stream = true;
    return RGBDFE_OK;
}

int rgbdfe_share_stream(rgbdfe_ctx* c, rgbdfe_ctx* owner)
{
    if (!c || !owner || c == owner || c->device != owner->device)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    rgbdfe_ctx* root = owner->stream_owner ? owner->stream_owner : owner; // sharing a sharer's stream = sharing its owner's
    if (root == c) {
        c->set_error("share_stream: the owner already enqueues on this context's stream");
        return RGBDFE_E_ARG;
    }
    release_stream(c);
    c->stream = root->stream;
    c->owns_stream = false;
    c->stream_owner = root;
    root->stream_sharers.push_back(c);
    return RGBDFE_OK;
}

int rgbdfe_set_option(rgbdfe_ctx* c, const char* name, int value)
{
    if (!c || !name)
        return RGBDFE_E_ARG;
    if (!strcmp(name, "depth_upload_of_8")) {
        if (value < 0 || value > 8)
            return RGBDFE_E_ARG;
        c->depth_upload_of_8 = value;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "depth_upload_phase")) {
        c->extract_calls = (unsigned)value;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "host_zero_copy")) {
        c->host_zero_copy = value != 0;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "overlap_blur")) {
        c->overlap_blur = value != 0;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "knn_waves")) {
        if (value < 1 || value > 64)
            return RGBDFE_E_ARG;
        c->knn_waves = value;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "chain_graphs")) {
        c->chain_graphs = value != 0;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "knn_dbg")) {
        c->knn_dbg = value;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "knn_umma")) {
        c->knn_umma = value < 0 ? 0 : (value > 2 ? 2 : value);
        return RGBDFE_OK;
    }
    if (!strcmp(name, "ransac_verify_cap")) {
        c->ransac_verify_cap = value;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "ransac_small_ctas")) {
        c->ransac_small_ctas = value != 0;
        return RGBDFE_OK;
    }
    if (!strcmp(name, "ransac_first_chunk")) {
        if (value < 1 || value > MAX_ITERS)
            return RGBDFE_E_ARG;
        c->ransac_first_chunk = value;
        return RGBDFE_OK;
    }
    c->set_error(std::string("unknown option ") + name);
    return RGBDFE_E_ARG;
}

int rgbdfe_sync(rgbdfe_ctx* c)
{
    if (!c)
        return RGBDFE_E_ARG;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return RGBDFE_OK;
}

int rgbdfe_orb_configure(rgbdfe_ctx* c, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    if (!c)
        return RGBDFE_E_ARG;
    if (nfeatures < 1 || nfeatures > c->max_feats || nlevels < 1 || nlevels > RGBDFE_MAX_LEVELS || !(scaleFactor > 1.0f) || iniThFAST < 0
        || iniThFAST > 255 || minThFAST < 0 || minThFAST > 255) {
        c->set_error("orb_configure: parameter out of range");
        return RGBDFE_E_ARG;
    }
    cudaSetDevice(c->device);
    c->nfeatures = nfeatures, c->scaleFactorF = scaleFactor, c->nlevels = nlevels, c->iniTh = iniThFAST, c->minTh = minThFAST;
    c->extractor = RGBDFE_EXTRACTOR_ORB2;
    // ORBextractor constructor tables (ORBextractor.cpp:348-383); scaleFactor is a double member
    const double sf = (double)scaleFactor;
    c->scale.assign(nlevels, 1.0f), c->sigma2.assign(nlevels, 1.0f), c->inv_scale.assign(nlevels, 1.0f), c->inv_sigma2.assign(nlevels, 1.0f);
    for (int i = 1; i < nlevels; i++) {
        c->scale[i] = (float)(c->scale[i - 1] * sf);
        c->sigma2[i] = c->scale[i] * c->scale[i];
    }
    for (int i = 0; i < nlevels; i++) {
        c->inv_scale[i] = 1.0f / c->scale[i];
        c->inv_sigma2[i] = 1.0f / c->sigma2[i];
    }
    c->quota.assign(nlevels, 0);
    float factor = (float)(1.0f / sf);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        c->quota[l] = cv_round_f(nDesired);
        sum += c->quota[l];
        nDesired *= factor;
    }
    c->quota[nlevels - 1] = std::max(nfeatures - sum, 0);

    // size the extraction workspaces for the largest image
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    free_orb_workspaces(c);
    size_t pyr = 0, blur = 0;
    int cells = 0, cand = 0, sel = 0, tx = 0, ty = 0;
    for (int l = 0; l < nlevels; l++) {
        const int w = cv_round_f((float)c->max_w * c->inv_scale[l]), h = cv_round_f((float)c->max_h * c->inv_scale[l]);
        pyr += (size_t)align_up(w + 38, 64) * (h + 38);
        blur += (size_t)align_up(w, 64) * h;
        const int nc = std::max((w - 32) / 30, 1), nr = std::max((h - 32) / 30, 1);
        cells += nc * nr;
        cand += nc * nr * CELL_CAP;
        sel += c->quota[l] + 16;
        tx += align_up(w + 38, 64), ty += (h + 38 + PYR_ROWS - 1) / PYR_ROWS;
    }
    const size_t B = (size_t)c->max_batch;
    c->img_slot = (size_t)align_up(c->max_w, 64) * c->max_h;
    c->depth_slot = (size_t)c->max_w * c->max_h;
    c->pyr_slot = align_up((int)pyr, 256), c->blur_slot = align_up((int)blur, 256);
    c->cell_slot = cells, c->cand_slot = cand, c->sel_slot = sel, c->tab_x_cap = tx, c->tab_y_cap = ty;
    CUDA_TRY(c, dmalloc(&c->d_img, B * c->img_slot + 16)); // + 16: k_pyramid_base reads whole aligned words
    CUDA_TRY(c, dmalloc(&c->d_depth, B * c->depth_slot));
    CUDA_TRY(c, dmalloc(&c->d_pyr, B * c->pyr_slot + 256));
    CUDA_TRY(c, dmalloc(&c->d_blur, B * c->blur_slot + 256));
    CUDA_TRY(c, dmalloc(&c->d_cell_count, B * cells));
    CUDA_TRY(c, dmalloc(&c->d_cell_list, B * cells * CELL_CAP));
    CUDA_TRY(c, dmalloc(&c->d_cand, B * cand));
    CUDA_TRY(c, dmalloc(&c->d_label, B * cand));
    CUDA_TRY(c, dmalloc(&c->d_cand_count, B * RGBDFE_MAX_LEVELS));
    CUDA_TRY(c, dmalloc(&c->d_sel, B * sel));
    CUDA_TRY(c, dmalloc(&c->d_sel_count, B * RGBDFE_MAX_LEVELS));
    CUDA_TRY(c, dmalloc(&c->d_tab_xpad, (size_t)tx));
    CUDA_TRY(c, dmalloc(&c->d_tab_group, (size_t)ty));
    CUDA_TRY(c, cudaMemset(c->d_sel_count, 0, B * RGBDFE_MAX_LEVELS * sizeof(int)));
    CUDA_TRY(c, cudaMemset(c->d_cand_count, 0, B * RGBDFE_MAX_LEVELS * sizeof(int)));
    memset(&c->cfg, 0, sizeof(c->cfg)); // geometry is rebuilt on the next extract
    c->configured = true;
    return RGBDFE_OK;
}

int rgbdfe_orb_tables(rgbdfe_ctx* c, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* quota)
{
    if (!c || !c->configured)
        return RGBDFE_E_STATE;
    for (int i = 0; i < c->nlevels; i++) {
        if (scale)
            scale[i] = c->scale[i];
        if (inv_scale)
            inv_scale[i] = c->inv_scale[i];
        if (sigma2)
            sigma2[i] = c->sigma2[i];
        if (inv_sigma2)
            inv_sigma2[i] = c->inv_sigma2[i];
        if (quota)
            quota[i] = c->quota[i];
    }
    return RGBDFE_OK;
}

} // extern "C"

namespace {

// 0: pageable / unknown host memory, 1: device (or managed) memory, 2: pinned host memory mapped into the device
// address space.  *dev receives the pointer kernels can dereference (kinds 1 and 2).
int pointer_kind(const void* p, const void** dev)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) {
        *dev = a.devicePointer ? a.devicePointer : p;
        return 1;
    }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) {
        *dev = a.devicePointer;
        return 2;
    }
    return 0;
}

// Kernel-visible pointers of a batch of images that are read where they are: device memory, or (depth image, option
// "host_zero_copy") pinned host memory, of which K6 then pulls only the sectors under the keypoints over PCIe
// instead of the whole map being uploaded.
bool in_place_pointers(const void* const* img, int count, bool pinned_host_too, const void** out)
{
    const void *d0 = nullptr, *d1 = nullptr;
    const int k0 = pointer_kind(img[0], &d0), k1 = pointer_kind(img[count - 1], &d1);
    if (k0 == 0 || k1 != k0 || (k0 == 2 && !pinned_host_too))
        return false;
    if (d0 == img[0] && d1 == img[count - 1]) { // unified addressing: the pointers are usable as they are
        for (int i = 0; i < count; i++)
            out[i] = img[i];
        return true;
    }
    for (int i = 0; i < count; i++)
        if (pointer_kind(img[i], &out[i]) != k0)
            return false;
    return true;
}

// The colour upload slots are three times the gray ones; they are allocated on first use.
int ensure_img_slots(rgbdfe_ctx* c, int channels)
{
    if (channels <= c->img_channels)
        return RGBDFE_OK;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    cudaFree(c->d_img);
    c->d_img = nullptr;
    CUDA_TRY(c, dmalloc(&c->d_img, (size_t)c->max_batch * c->img_slot * channels + 16));
    c->img_channels = channels;
    return RGBDFE_OK;
}

// cv::undistortPoints(p, K, distCoef, noArray(), K) for one point, in double as OpenCV (undistort.dispatch.cpp); the
// device runs the same recurrence for the keypoints, the host needs it for the four image corners only.
void undistort_point(const FramePrep& P, float u, float v, float* ou, float* ov)
{
    double x = (u - P.cx) * P.ifx, y = (v - P.cy) * P.ify;
    const double x0 = x, y0 = y;
    for (int it = 0; it < 5; it++) {
        const double r2 = x * x + y * y;
        const double icdist = 1.0 / (1.0 + ((P.k[4] * r2 + P.k[1]) * r2 + P.k[0]) * r2);
        if (icdist < 0) {
            x = x0, y = y0;
            break;
        }
        const double dx = 2.0 * P.k[2] * x * y + P.k[3] * (r2 + 2.0 * x * x);
        const double dy = P.k[2] * (r2 + 2.0 * y * y) + 2.0 * P.k[3] * x * y;
        x = (x0 - dx) * icdist;
        y = (y0 - dy) * icdist;
    }
    *ou = (float)(P.fx * x + P.cx);
    *ov = (float)(P.fy * y + P.cy);
}

// Frame preprocessing constants: intrinsics, distortion, image bounds (Frame::computeImageBounds, Core/Frame.cpp:283-316)
// and grid cell sizes (:65-66).
FramePrep make_frame_prep(int w, int h, float fx, float fy, float cx, float cy, const float* dist, float depth16_factor)
{
    FramePrep P;
    memset(&P, 0, sizeof(P));
    P.fx = fx, P.fy = fy, P.cx = cx, P.cy = cy;
    P.ifx = fx != 0.0f ? 1.0 / P.fx : 0.0, P.ify = fy != 0.0f ? 1.0 / P.fy : 0.0;
    P.cxf = cx, P.cyf = cy;
    P.invfx = fx != 0.0f ? 1.0f / fx : 0.0f, P.invfy = fy != 0.0f ? 1.0f / fy : 0.0f;
    P.depth16_factor = depth16_factor;
    P.undistort = dist && dist[0] != 0.0f;
    float minX = 0.0f, maxX = (float)w, minY = 0.0f, maxY = (float)h;
    if (P.undistort) {
        for (int i = 0; i < 5; i++)
            P.k[i] = dist[i];
        float x[4], y[4];
        undistort_point(P, 0.0f, 0.0f, &x[0], &y[0]);
        undistort_point(P, (float)w, 0.0f, &x[1], &y[1]);
        undistort_point(P, 0.0f, (float)h, &x[2], &y[2]);
        undistort_point(P, (float)w, (float)h, &x[3], &y[3]);
        minX = std::min(x[0], x[2]), maxX = std::max(x[1], x[3]);
        minY = std::min(y[0], y[1]), maxY = std::max(y[2], y[3]);
    }
    P.minX = minX, P.minY = minY, P.maxX = maxX, P.maxY = maxY;
    P.gridWInv = 64.0f / (maxX - minX); // FRAME_GRID_COLS / (mnMaxX - mnMinX)
    P.gridHInv = 48.0f / (maxY - minY);
    return P;
}

// Host images -> the context's upload slots; out[i] = where image i landed.  Images with dense rows that follow
// each other in the caller's memory (a frame ring buffer, ascending or descending) go up as ONE strided copy per
// run: a 64-frame batch is otherwise bound by the fixed cost of its 128 separate transfers, not by PCIe.
int upload_images(rgbdfe_ctx* c, const void* const* img, int count, size_t row_bytes, int rows, size_t src_pitch, uint8_t* slots, size_t slot_bytes,
    const void** out)
{
    const size_t bytes = row_bytes * rows;
    int i = 0;
    while (i < count) {
        int j = i, dir = 0;
        while (src_pitch == row_bytes && j + 1 < count) {
            const ptrdiff_t d = static_cast<const char*>(img[j + 1]) - static_cast<const char*>(img[j]);
            const int nd = d == (ptrdiff_t)bytes ? 1 : (d == -(ptrdiff_t)bytes ? -1 : 0);
            if (nd == 0 || (dir != 0 && nd != dir))
                break;
            dir = nd;
            j++;
        }
        if (j == i)
            CUDA_TRY(c, cudaMemcpy2DAsync(slots + i * slot_bytes, row_bytes, img[i], src_pitch, row_bytes, rows, cudaMemcpyDefault, c->stream));
        else // one "row" per image: source pitch = image size, destination pitch = slot size
            CUDA_TRY(c, cudaMemcpy2DAsync(slots + i * slot_bytes, slot_bytes, dir > 0 ? img[i] : img[j], bytes, bytes, j - i + 1, cudaMemcpyDefault, c->stream));
        for (int m = i; m <= j; m++)
            out[m] = slots + (size_t)(dir >= 0 ? m : i + (j - m)) * slot_bytes;
        i = j + 1;
    }
    return RGBDFE_OK;
}

// Upload `count` gray (+ optional depth) images and run K1..K6 on them as one batch.
// channels: 1 (gray) or 3 (BGR, converted on the way into level 0); dist: k1 k2 p1 p2 k3 or nullptr.
int run_extract(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, const uint8_t* const* gray, int w, int h, int stride,
    const float* const* depth, int depth_stride, float fx, float fy, float cx, float cy, const uint16_t* const* depth16 = nullptr,
    float depth16_factor = 0.0f, int channels = 1, const float* dist = nullptr)
{
    const bool cvorb = c->extractor == RGBDFE_EXTRACTOR_ORB; // Extractor(ORB, ORB): the cv::ORB branch (orb_cv.cu)
    if (!c->configured)
        return RGBDFE_E_STATE;
    if (count < 1 || count > c->max_batch || !gray || !frames) {
        c->set_error("extract: count outside [1, max_batch]");
        return RGBDFE_E_ARG;
    }
    if (cvorb && channels != 1) {
        c->set_error("extract: the cv::ORB branch takes gray images");
        return RGBDFE_E_ARG;
    }
    cudaSetDevice(c->device);
    int rc = cvorb ? RGBDFE_OK : build_level_config(c, w, h);
    if (rc != RGBDFE_OK)
        return rc;
    FrameBatch fb;
    ImgPtrs gp, dp;
    memset(&fb, 0, sizeof(fb));
    memset(&gp, 0, sizeof(gp));
    memset(&dp, 0, sizeof(dp));
    const void* const* dany = depth ? (const void* const*)depth : (const void* const*)depth16;
    const size_t delem = depth ? 4 : 2;
    for (int i = 0; i < count; i++)
        if (!gray[i] || !frames[i] || (dany && !dany[i]))
            return RGBDFE_E_ARG;
    // Images in device memory are read in place (all pointers of one call must be of one kind; the first and last
    // are checked); host images are uploaded into the context's slots, except a pinned depth map with "host_zero_copy".
    if (channels != 1 && channels != 3)
        return RGBDFE_E_ARG;
    if ((rc = ensure_img_slots(c, channels)) != RGBDFE_OK)
        return rc;
    const bool gray_in_place = in_place_pointers((const void* const*)gray, count, false, gp.p);
    // "depth_upload_of_8": of every 8 calls, that many upload a pinned depth map whole (copy engine) and the others read it
    // in place (PCIe sector reads): the two transports have different limits and a mix uses both
    const bool zero_copy_now = c->host_zero_copy && (int)(c->extract_calls++ & 7u) >= c->depth_upload_of_8;
    const bool depth_in_place = dany && in_place_pointers(dany, count, zero_copy_now, dp.p);
    int gstride = stride, dstride = depth_stride;
    for (int i = 0; i < count; i++)
        fb.f[i] = *frames[i];
    if (!gray_in_place) {
        if ((rc = upload_images(c, (const void* const*)gray, count, (size_t)w * channels, h, (size_t)stride, reinterpret_cast<uint8_t*>(c->d_img),
                 c->img_slot * c->img_channels, gp.p))
            != RGBDFE_OK)
            return rc;
        gstride = w * channels;
    }
    if (dany && !depth_in_place) {
        if ((rc = upload_images(c, dany, count, (size_t)w * delem, h, (size_t)depth_stride * delem, reinterpret_cast<uint8_t*>(c->d_depth),
                 c->depth_slot * sizeof(float), dp.p))
            != RGBDFE_OK)
            return rc;
        dstride = w;
    }
    if (cvorb)
        return launch_cvorb(c, count, w, h, gp, gstride, fb, dany != nullptr, dp, dstride,
            make_frame_prep(w, h, fx, fy, cx, cy, dist, depth16 ? depth16_factor : 0.0f));
    if ((rc = launch_pyramid(c, count, gp, gstride, channels)) != RGBDFE_OK)
        return rc;
    if ((rc = launch_fast(c, count)) != RGBDFE_OK)
        return rc;
    if (c->overlap_blur && !c->profiling) {
        // fork after K2: the blur on the side stream, the quadtree on the context's stream, joined before K6
        if (!c->side_stream) {
            CUDA_TRY(c, cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking));
            CUDA_TRY(c, cudaEventCreateWithFlags(&c->side_fork, cudaEventDisableTiming));
            CUDA_TRY(c, cudaEventCreateWithFlags(&c->side_join, cudaEventDisableTiming));
        }
        CUDA_TRY(c, cudaEventRecord(c->side_fork, c->stream));
        CUDA_TRY(c, cudaStreamWaitEvent(c->side_stream, c->side_fork, 0));
        std::swap(c->stream, c->side_stream);
        rc = launch_blur(c, count);
        std::swap(c->stream, c->side_stream);
        if (rc != RGBDFE_OK)
            return rc;
        CUDA_TRY(c, cudaEventRecord(c->side_join, c->side_stream));
        if ((rc = launch_quadtree(c, count)) != RGBDFE_OK)
            return rc;
        CUDA_TRY(c, cudaStreamWaitEvent(c->stream, c->side_join, 0));
    } else {
        if ((rc = launch_blur(c, count)) != RGBDFE_OK)
            return rc;
        if ((rc = launch_quadtree(c, count)) != RGBDFE_OK)
            return rc;
    }
    const void* unused = nullptr;
    const bool depth_on_host = depth_in_place && pointer_kind(dany[0], &unused) == 2; // pinned host memory read over PCIe
    if ((rc = launch_describe(c, count, fb, dany != nullptr, dp, dstride, make_frame_prep(w, h, fx, fy, cx, cy, dist, depth16 ? depth16_factor : 0.0f),
             depth_on_host))
        != RGBDFE_OK)
        return rc;
    return RGBDFE_OK;
}

int set_pair(rgbdfe_ctx* c, int slot, const MatchPair& p)
{
    k_set_pair<<<1, 1, 0, c->stream>>>(c->d_pairs + slot, p);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

Se3Params make_params(rgbdfe_ctx* c, int iterations, unsigned min_inl, float max_mahal, unsigned sample)
{
    Se3Params p;
    p.iterations = iterations;
    p.min_inlier_th = min_inl;
    p.max_mahal = max_mahal;
    p.sample_size = sample;
    p.rcx = c->rcx;
    p.rcy = c->rcy;
    return p;
}

} // namespace

extern "C" {

int rgbdfe_orb_extract(rgbdfe_ctx* c, const uint8_t* gray, int w, int h, int stride, rgbdfe_keypoint* kps, uint8_t* desc, int cap, int* n)
{
    if (!c || !n)
        return RGBDFE_E_ARG;
    *n = 0;
    if (!gray || w <= 0 || h <= 0)
        return RGBDFE_OK; // ORBextractor::operator() returns silently on an empty image (ORBextractor.cpp:708-709)
    rgbdfe_frame* f = c->scratch;
    int rc = run_extract(c, &f, 1, &gray, w, h, stride, nullptr, 0, 0, 0, 0, 0);
    if (rc != RGBDFE_OK)
        return rc;
    return rgbdfe_frame_download(c, f, kps, desc, nullptr, nullptr, cap, n);
}

int rgbdfe_cvorb_configure(rgbdfe_ctx* c, int nfeatures, float scaleFactor, int nlevels)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    const int rc = cvorb_configure(c, nfeatures, scaleFactor, nlevels);
    if (rc == RGBDFE_OK) {
        c->extractor = RGBDFE_EXTRACTOR_ORB;
        c->configured = true;
    }
    return rc;
}

int rgbdfe_cvorb_stage_level(rgbdfe_ctx* c, int level, int which, uint8_t* dst, int dst_stride, int* w, int* h)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    return cvorb_stage_level(c, level, which, dst, dst_stride, w, h);
}

int rgbdfe_cvorb_stage_harris(rgbdfe_ctx* c, int level, rgbdfe_keypoint* out, int cap, int* n)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    return cvorb_stage_harris(c, level, out, cap, n);
}

int rgbdfe_orb_pyramid_level(rgbdfe_ctx* c, int level, uint8_t* dst, int dst_stride, int* w, int* h)
{
    if (!c || level < 0 || level >= c->cfg.nlevels || c->cfg.w == 0)
        return RGBDFE_E_STATE;
    const LevelDesc& L = c->cfg.lv[level];
    if (w)
        *w = L.w;
    if (h)
        *h = L.h;
    if (dst) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        CUDA_TRY(c, cudaMemcpy2D(dst, dst_stride, c->d_pyr + L.pyr_off, L.pstride, L.w + 38, L.h + 38, cudaMemcpyDeviceToHost));
    }
    return RGBDFE_OK;
}

int rgbdfe_orb_stage_blurred(rgbdfe_ctx* c, int level, uint8_t* dst, int dst_stride)
{
    if (!c || level < 0 || level >= c->cfg.nlevels || c->cfg.w == 0 || !dst)
        return RGBDFE_E_STATE;
    const LevelDesc& L = c->cfg.lv[level];
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy2D(dst, dst_stride, c->d_blur + L.blur_off, L.bstride, L.w, L.h, cudaMemcpyDeviceToHost));
    return RGBDFE_OK;
}

int rgbdfe_orb_stage_candidates(rgbdfe_ctx* c, int level, rgbdfe_keypoint* out, int cap, int* n)
{
    if (!c || level < 0 || level >= c->cfg.nlevels || c->cfg.w == 0 || !n)
        return RGBDFE_E_STATE;
    const LevelDesc& L = c->cfg.lv[level];
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int cnt = 0;
    CUDA_TRY(c, cudaMemcpy(&cnt, c->d_cand_count + level, sizeof(int), cudaMemcpyDeviceToHost));
    *n = cnt;
    const int m = std::min(cnt, cap);
    if (m > 0 && out) {
        std::vector<unsigned> tmp(m);
        CUDA_TRY(c, cudaMemcpy(tmp.data(), c->d_cand + L.cand_off, (size_t)m * 4, cudaMemcpyDeviceToHost));
        for (int i = 0; i < m; i++) {
            out[i].x = (float)cand_x(tmp[i]);
            out[i].y = (float)cand_y(tmp[i]);
            out[i].size = 7.f;
            out[i].angle = -1.f;
            out[i].response = (float)cand_s(tmp[i]);
            out[i].octave = 0;
            out[i].class_id = -1;
        }
    }
    return RGBDFE_OK;
}

int rgbdfe_camera_bounds(const rgbdfe_camera* cam, int w, int h, float bounds[4], float grid_inv[2])
{
    if (!cam || !bounds || !grid_inv || w <= 0 || h <= 0)
        return RGBDFE_E_ARG;
    const FramePrep P = make_frame_prep(w, h, cam->fx, cam->fy, cam->cx, cam->cy, cam->dist, 0.0f);
    bounds[0] = P.minX, bounds[1] = P.maxX, bounds[2] = P.minY, bounds[3] = P.maxY;
    grid_inv[0] = P.gridWInv, grid_inv[1] = P.gridHInv;
    return RGBDFE_OK;
}

int rgbdfe_frame_create(rgbdfe_ctx* c, rgbdfe_frame** frame)
{
    if (!c || !frame)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    return alloc_frame(c, frame);
}

void rgbdfe_frame_destroy(rgbdfe_ctx* c, rgbdfe_frame* frame)
{
    if (c) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
    }
    free_frame(frame);
}

int rgbdfe_frames_extract(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, const uint8_t* const* gray, int w, int h, int stride,
    const float* const* depth, int depth_stride_elems, float fx, float fy, float cx, float cy)
{
    if (!c)
        return RGBDFE_E_ARG;
    return run_extract(c, frames, count, gray, w, h, stride, depth, depth_stride_elems, fx, fy, cx, cy);
}

int rgbdfe_frames_extract_raw(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, const uint8_t* const* gray, int w, int h, int stride,
    const uint16_t* const* depth16, int depth_stride_elems, float depth_map_factor, float fx, float fy, float cx, float cy)
{
    if (!c || !depth16 || !(depth_map_factor > 0.0f))
        return RGBDFE_E_ARG;
    return run_extract(c, frames, count, gray, w, h, stride, nullptr, depth_stride_elems, fx, fy, cx, cy, depth16, depth_map_factor);
}

int rgbdfe_frames_from_images(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, const uint8_t* const* color, int channels, int w, int h,
    int stride, const uint16_t* const* depth16, int depth_stride_elems, const rgbdfe_camera* cam)
{
    if (!c || !cam || !depth16 || !(cam->depth_map_factor > 0.0f))
        return RGBDFE_E_ARG;
    return run_extract(c, frames, count, color, w, h, stride, nullptr, depth_stride_elems, cam->fx, cam->fy, cam->cx, cam->cy, depth16,
        cam->depth_map_factor, channels, cam->dist);
}

int rgbdfe_frame_download_undistorted(rgbdfe_ctx* c, rgbdfe_frame* f, float* pts_un, int* grid_cell, int cap, int* n)
{
    if (!c || !f || !n)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int cnt = 0;
    CUDA_TRY(c, cudaMemcpy(&cnt, f->n, sizeof(int), cudaMemcpyDeviceToHost));
    *n = cnt;
    if (cnt < 0)
        return RGBDFE_E_CAPACITY;
    const int m = std::min(cnt, cap);
    if (m > 0 && pts_un)
        CUDA_TRY(c, cudaMemcpy(pts_un, f->pt_un, (size_t)m * sizeof(float2), cudaMemcpyDeviceToHost));
    if (m > 0 && grid_cell)
        CUDA_TRY(c, cudaMemcpy(grid_cell, f->cell, (size_t)m * sizeof(int), cudaMemcpyDeviceToHost));
    return RGBDFE_OK;
}

int rgbdfe_frame_download(rgbdfe_ctx* c, rgbdfe_frame* f, rgbdfe_keypoint* kps, uint8_t* desc, rgbdfe_point3f* xyz, uint8_t* outlier, int cap,
    int* n)
{
    if (!c || !f || !n)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int cnt = 0;
    CUDA_TRY(c, cudaMemcpy(&cnt, f->n, sizeof(int), cudaMemcpyDeviceToHost));
    *n = cnt;
    if (cnt < 0) {
        c->set_error("extractor capacity overflow (quadtree nodes, or more than 65535 / w*h/8 FAST corners on one level of the cv::ORB branch)");
        return RGBDFE_E_CAPACITY;
    }
    const int m = std::min(cnt, cap);
    if (m > 0) {
        if (kps)
            CUDA_TRY(c, cudaMemcpy(kps, f->kps, (size_t)m * sizeof(rgbdfe_keypoint), cudaMemcpyDeviceToHost));
        if (desc)
            CUDA_TRY(c, cudaMemcpy(desc, f->desc, (size_t)m * 32, cudaMemcpyDeviceToHost));
        if (xyz)
            CUDA_TRY(c, cudaMemcpy(xyz, f->xyz, (size_t)m * sizeof(rgbdfe_point3f), cudaMemcpyDeviceToHost));
        if (outlier)
            CUDA_TRY(c, cudaMemcpy(outlier, f->outlier, (size_t)m, cudaMemcpyDeviceToHost));
    }
    return cnt > cap ? RGBDFE_E_CAPACITY : RGBDFE_OK;
}

int rgbdfe_frames_download_async(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, rgbdfe_keypoint* kps, uint8_t* desc, rgbdfe_point3f* xyz, int cap,
    int* n)
{
    if (!c || !frames || count < 0 || count > 64 || cap < 1 || !n)
        return RGBDFE_E_ARG;
    if (count == 0)
        return RGBDFE_OK;
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    FrameBatch fb;
    for (int i = 0; i < count; i++) {
        if (!frames[i])
            return RGBDFE_E_ARG;
        fb.f[i] = *frames[i];
    }
    // The frames are separate allocations: one kernel packs their valid rows into a staging block laid out like the
    // outputs, and each output array then leaves in ONE copy (a batch of 64 frames would otherwise be 256 small ones).
    const size_t row_words = 7 + 8 + 3; // KeyPoint 28 B, descriptor 32 B, Point3f 12 B
    const size_t need = (size_t)count * cap * row_words * 4 + (size_t)count * 4;
    if (need > c->dl_bytes) {
        CUDA_TRY(c, cudaStreamSynchronize(s));
        cudaFree(c->d_dl);
        c->d_dl = nullptr, c->dl_bytes = 0;
        CUDA_TRY(c, cudaMalloc((void**)&c->d_dl, need));
        c->dl_bytes = need;
    }
    unsigned* d_kps = reinterpret_cast<unsigned*>(c->d_dl);
    unsigned* d_desc = d_kps + (size_t)count * cap * 7;
    unsigned* d_xyz = d_desc + (size_t)count * cap * 8;
    int* d_n = reinterpret_cast<int*>(d_xyz + (size_t)count * cap * 3);
    k_pack_frames<<<dim3(8, count), 256, 0, s>>>(fb, kps ? d_kps : nullptr, desc ? d_desc : nullptr, xyz ? d_xyz : nullptr, d_n, cap);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(n, d_n, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (kps)
        CUDA_TRY(c, cudaMemcpyAsync(kps, d_kps, (size_t)count * cap * 28, cudaMemcpyDeviceToHost, s));
    if (desc)
        CUDA_TRY(c, cudaMemcpyAsync(desc, d_desc, (size_t)count * cap * 32, cudaMemcpyDeviceToHost, s));
    if (xyz)
        CUDA_TRY(c, cudaMemcpyAsync(xyz, d_xyz, (size_t)count * cap * 12, cudaMemcpyDeviceToHost, s));
    return RGBDFE_OK;
}

int rgbdfe_frame_upload(rgbdfe_ctx* c, rgbdfe_frame* f, const uint8_t* desc, const rgbdfe_point3f* xyz, int n)
{
    if (!c || !f || n < 0 || (n > 0 && (!desc || !xyz)))
        return RGBDFE_E_ARG;
    if (n > f->cap) {
        c->set_error("frame_upload: descriptor count exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    if (n > 0) {
        CUDA_TRY(c, cudaMemcpyAsync(f->desc, desc, (size_t)n * 32, cudaMemcpyHostToDevice, s));
        CUDA_TRY(c, cudaMemcpyAsync(f->xyz, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, s));
    }
    k_set_int<<<1, 1, 0, s>>>(f->n, n);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

int rgbdfe_frames_match(rgbdfe_ctx* c, rgbdfe_frame* Q, rgbdfe_frame* T, int nq, int nt, const uint8_t* q_outlier, float ratio, int discard_outliers,
    rgbdfe_dmatch* out, int cap, int* n)
{
    if (!c || !Q || !T || !n || nq < 0 || nt < 0 || nq > Q->cap || nt > T->cap)
        return RGBDFE_E_ARG;
    *n = 0;
    if (nq == 0 || nt == 0)
        return RGBDFE_OK;
    cudaSetDevice(c->device);
    int rc = ensure_pairs(c, 1);
    if (rc != RGBDFE_OK)
        return rc;
    cudaStream_t s = c->stream;
    if (q_outlier)
        CUDA_TRY(c, cudaMemcpyAsync(Q->outlier, q_outlier, (size_t)nq, cudaMemcpyHostToDevice, s));
    MatchPair p = { Q->desc, T->desc, Q->n, T->n, q_outlier ? Q->outlier : nullptr, Q->xyz, T->xyz, nullptr };
    if ((rc = set_pair(c, 0, p)) != RGBDFE_OK)
        return rc;
    if ((rc = launch_match(c, c->d_pairs, 1, nq, nt, ratio, discard_outliers, nullptr)) != RGBDFE_OK)
        return rc;
    // count and matches leave together through the pinned result staging: one synchronisation
    int* h_cnt = reinterpret_cast<int*>(c->h_results);
    rgbdfe_dmatch* h_m = reinterpret_cast<rgbdfe_dmatch*>(h_cnt + 4);
    const int room = (int)std::min<size_t>((size_t)std::min(cap, nq), (sizeof(rgbdfe_se3_result) * c->result_cap - 16) / sizeof(rgbdfe_dmatch));
    CUDA_TRY(c, cudaMemcpyAsync(h_cnt, c->d_nmatch, sizeof(int), cudaMemcpyDeviceToHost, s));
    if (room > 0 && out)
        CUDA_TRY(c, cudaMemcpyAsync(h_m, c->d_matches, (size_t)room * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    const int cnt = *h_cnt;
    *n = cnt;
    const int m = std::min(cnt, cap);
    if (m > 0 && out) {
        if (m <= room)
            memcpy(out, h_m, (size_t)m * sizeof(rgbdfe_dmatch));
        else
            CUDA_TRY(c, cudaMemcpy(out, c->d_matches, (size_t)m * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost));
    }
    return cnt > cap ? RGBDFE_E_CAPACITY : RGBDFE_OK;
}

int rgbdfe_match_knn2(rgbdfe_ctx* c, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt, const uint8_t* q_outlier,
    const rgbdfe_point3f* q_xyz, const rgbdfe_point3f* t_xyz, float ratio, int discard_outliers, rgbdfe_dmatch* out, int cap, int* n)
{
    if (!c || !n || nq < 0 || nt < 0)
        return RGBDFE_E_ARG;
    *n = 0;
    if (nq > c->kp_cap || nt > c->kp_cap) {
        c->set_error("match: descriptor count exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    if (nq == 0 || nt == 0)
        return RGBDFE_OK;
    if (!q_desc || !t_desc || !q_xyz || !t_xyz)
        return RGBDFE_E_ARG;
    int rc = rgbdfe_frame_upload(c, c->scratch, q_desc, q_xyz, nq);
    if (rc == RGBDFE_OK)
        rc = rgbdfe_frame_upload(c, c->scratch2, t_desc, t_xyz, nt);
    if (rc != RGBDFE_OK)
        return rc;
    return rgbdfe_frames_match(c, c->scratch, c->scratch2, nq, nt, q_outlier, ratio, discard_outliers, out, cap, n);
}

int rgbdfe_rng_seed(rgbdfe_ctx* c, unsigned seed)
{
    if (!c)
        return RGBDFE_E_ARG;
    // glibc srandom_r (TYPE_3): 31-word LCG fill, then 310 discarded outputs
    RngState s;
    unsigned lin[344];
    if (seed == 0)
        seed = 1;
    lin[0] = seed;
    for (int i = 1; i < 31; i++) {
        long long w = (16807LL * (int)lin[i - 1]) % 2147483647;
        if (w < 0)
            w += 2147483647;
        lin[i] = (unsigned)w;
    }
    for (int i = 31; i < 34; i++)
        lin[i] = lin[i - 31];
    for (int i = 34; i < 344; i++)
        lin[i] = lin[i - 31] + lin[i - 3];
    for (int i = 344 - 34; i < 344; i++)
        s.r[i % 34] = (int)lin[i];
    s.idx = 344;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(c->d_rng, &s, sizeof(s), cudaMemcpyHostToDevice));
    return RGBDFE_OK;
}

int rgbdfe_rng_get_state(rgbdfe_ctx* c, int32_t* state35)
{
    if (!c || !state35)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(state35, c->d_rng, sizeof(RngState), cudaMemcpyDeviceToHost));
    return RGBDFE_OK;
}

int rgbdfe_rng_set_state(rgbdfe_ctx* c, const int32_t* state35)
{
    if (!c || !state35)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(c->d_rng, state35, sizeof(RngState), cudaMemcpyHostToDevice));
    return RGBDFE_OK;
}

int rgbdfe_rng_next(rgbdfe_ctx* c, int* value)
{
    if (!c || !value)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    RngState s;
    CUDA_TRY(c, cudaMemcpy(&s, c->d_rng, sizeof(s), cudaMemcpyDeviceToHost));
    // rand(): r[i] = r[i-3] + r[i-31], output >> 1 (glibc TYPE_3)
    const unsigned v = (unsigned)s.r[(s.idx - 31) % 34] + (unsigned)s.r[(s.idx - 3) % 34];
    s.r[s.idx % 34] = (int)v;
    s.idx++;
    if (s.idx >= 34 * 1000000)
        s.idx -= 34 * 999990;
    *value = (int)(v >> 1);
    CUDA_TRY(c, cudaMemcpy(c->d_rng, &s, sizeof(s), cudaMemcpyHostToDevice));
    return RGBDFE_OK;
}

int rgbdfe_depth_cov_get(rgbdfe_ctx* c, int* latched, double* cov)
{
    if (!c)
        return RGBDFE_E_ARG;
    DepthLatch l;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(&l, c->d_latch, sizeof(l), cudaMemcpyDeviceToHost));
    if (latched)
        *latched = l.latched;
    if (cov)
        *cov = l.cov;
    return RGBDFE_OK;
}

int rgbdfe_depth_cov_set(rgbdfe_ctx* c, int latched, double cov)
{
    if (!c)
        return RGBDFE_E_ARG;
    DepthLatch l = { latched, 0, cov };
    c->latch_known = latched != 0;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(c->d_latch, &l, sizeof(l), cudaMemcpyHostToDevice));
    return RGBDFE_OK;
}

} // extern "C"

extern "C" {

int rgbdfe_frames_ransac_se3(rgbdfe_ctx* c, rgbdfe_frame* F1, rgbdfe_frame* F2, int n1, int n2, const rgbdfe_dmatch* matches, int nm, int iterations,
    unsigned min_inlier_th, float max_mahal, unsigned sample_size, rgbdfe_se3_result* result, rgbdfe_dmatch* inliers, int cap)
{
    if (!c || !F1 || !F2 || !result || nm < 0 || n1 < 0 || n2 < 0 || (nm > 0 && !matches))
        return RGBDFE_E_ARG;
    if (n1 > c->kp_cap || n2 > c->kp_cap || nm > c->kp_cap) {
        c->set_error("ransac_se3: input exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    for (int i = 0; i < nm; i++)
        if (matches[i].queryIdx < 0 || matches[i].queryIdx >= n1 || matches[i].trainIdx < 0 || matches[i].trainIdx >= n2) {
            c->set_error("ransac_se3: match index out of range");
            return RGBDFE_E_ARG;
        }
    cudaSetDevice(c->device);
    int rc = ensure_pairs(c, 1);
    if (rc != RGBDFE_OK)
        return rc;
    cudaStream_t s = c->stream;
    if (nm > 0)
        CUDA_TRY(c, cudaMemcpyAsync(c->d_matches, matches, (size_t)nm * sizeof(rgbdfe_dmatch), cudaMemcpyHostToDevice, s));
    k_set_int<<<1, 1, 0, s>>>(c->d_nmatch, nm);
    c->launches++;
    MatchPair p = { nullptr, nullptr, F1->n, F2->n, nullptr, F1->xyz, F2->xyz, nullptr };
    if ((rc = set_pair(c, 0, p)) != RGBDFE_OK)
        return rc;
    Se3Params prm = make_params(c, iterations, min_inlier_th, max_mahal, sample_size);
    if ((rc = launch_ransac_se3(c, c->d_pairs, 0, 1, prm, 0, 0, 0, 0, 0, c->d_results, nullptr)) != RGBDFE_OK)
        return rc;
    // result, sorted match list and inlier positions leave together through pinned staging: one synchronisation
    const size_t stage_bytes = sizeof(rgbdfe_se3_result) * c->result_cap;
    unsigned char* hb = reinterpret_cast<unsigned char*>(c->h_results);
    const size_t off_sorted = 128, off_pos = off_sorted + (size_t)nm * sizeof(rgbdfe_dmatch);
    const bool staged = inliers && nm > 0 && off_pos + (size_t)nm * sizeof(int) <= stage_bytes;
    CUDA_TRY(c, cudaMemcpyAsync(hb, c->d_results, sizeof(*result), cudaMemcpyDeviceToHost, s));
    if (staged) {
        CUDA_TRY(c, cudaMemcpyAsync(hb + off_sorted, c->d_sorted, (size_t)nm * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(c, cudaMemcpyAsync(hb + off_pos, c->d_inl_pos, (size_t)nm * sizeof(int), cudaMemcpyDeviceToHost, s));
    }
    CUDA_TRY(c, cudaStreamSynchronize(s));
    memcpy(result, hb, sizeof(*result));
    if (inliers && result->n_inliers > 0) {
        if (staged) {
            const rgbdfe_dmatch* sorted = reinterpret_cast<const rgbdfe_dmatch*>(hb + off_sorted);
            const int* pos = reinterpret_cast<const int*>(hb + off_pos);
            for (int i = 0; i < result->n_inliers && i < cap; i++)
                inliers[i] = sorted[pos[i]];
        } else {
            std::vector<rgbdfe_dmatch> sorted(nm);
            std::vector<int> pos(result->n_inliers);
            CUDA_TRY(c, cudaMemcpy(sorted.data(), c->d_sorted, (size_t)nm * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost));
            CUDA_TRY(c, cudaMemcpy(pos.data(), c->d_inl_pos, (size_t)result->n_inliers * sizeof(int), cudaMemcpyDeviceToHost));
            for (int i = 0; i < result->n_inliers && i < cap; i++)
                inliers[i] = sorted[pos[i]];
        }
        if (result->n_inliers > cap)
            return RGBDFE_E_CAPACITY;
    }
    return RGBDFE_OK;
}

int rgbdfe_ransac_se3(rgbdfe_ctx* c, const rgbdfe_point3f* p1, int n1, const rgbdfe_point3f* p2, int n2, const rgbdfe_dmatch* matches, int nm,
    int iterations, unsigned min_inlier_th, float max_mahal, unsigned sample_size, rgbdfe_se3_result* result, rgbdfe_dmatch* inliers, int cap)
{
    if (!c || !result || nm < 0 || n1 < 0 || n2 < 0 || (nm > 0 && (!p1 || !p2 || !matches)))
        return RGBDFE_E_ARG;
    if (n1 > c->kp_cap || n2 > c->kp_cap || nm > c->kp_cap) {
        c->set_error("ransac_se3: input exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    cudaSetDevice(c->device);
    rgbdfe_frame *Q = c->scratch, *T = c->scratch2;
    cudaStream_t s = c->stream;
    if (n1 > 0)
        CUDA_TRY(c, cudaMemcpyAsync(Q->xyz, p1, (size_t)n1 * 12, cudaMemcpyHostToDevice, s));
    if (n2 > 0)
        CUDA_TRY(c, cudaMemcpyAsync(T->xyz, p2, (size_t)n2 * 12, cudaMemcpyHostToDevice, s));
    return rgbdfe_frames_ransac_se3(c, Q, T, n1, n2, matches, nm, iterations, min_inlier_th, max_mahal, sample_size, result, inliers, cap);
}

int rgbdfe_ransac_umeyama(rgbdfe_ctx* c, const rgbdfe_point3f* p1, int n1, const rgbdfe_point3f* p2, int n2, const rgbdfe_dmatch* matches, int nm,
    const rgbdfe_umeyama_params* params, rgbdfe_umeyama_result* result, rgbdfe_dmatch* inliers, int cap)
{
    if (!c || !result || !params || nm < 0 || n1 < 0 || n2 < 0 || (nm > 0 && (!p1 || !p2 || !matches)))
        return RGBDFE_E_ARG;
    if (params->errorVersion != 0 && params->errorVersion != 4) {
        // REPROJECTION / MAHALANOBIS versions dereference a null camera or read an uninitialised covariance
        // in the reference (Ransac.cpp:15,282-321,350); they are not executable there and are refused here
        c->set_error("ransac_umeyama: only EUCLIDEAN_ERROR (0) and ADAPTIVE_ERROR (4) are executable in the reference");
        return RGBDFE_E_ARG;
    }
    if (params->usedPairs < 1 || params->usedPairs > 8) {
        c->set_error("ransac_umeyama: usedPairs must be in [1, 8]");
        return RGBDFE_E_ARG;
    }
    if (n1 > c->kp_cap || n2 > c->kp_cap || nm > c->kp_cap) {
        c->set_error("ransac_umeyama: input exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    int M = 0; // matches that pass the depth gate (Ransac.cpp:72-83); only the count is needed on the host
    for (int i = 0; i < nm; i++) {
        const rgbdfe_dmatch& m = matches[i];
        if (m.queryIdx < 0 || m.queryIdx >= n1 || m.trainIdx < 0 || m.trainIdx >= n2) {
            c->set_error("ransac_umeyama: match index out of range");
            return RGBDFE_E_ARG;
        }
        const float z1 = p1[m.queryIdx].z, z2 = p2[m.trainIdx].z;
        if (!(z1 < 0.1f || z1 > 6.0f || z2 < 0.1f || z2 > 6.0f))
            M++;
    }
    result->ok = 0;
    for (int i = 0; i < 16; i++)
        result->T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    result->n_inliers = 0;
    result->iters_executed = 0;
    if (M < params->minimalNumberOfMatches) // Ransac.cpp:86-89
        return RGBDFE_OK;
    if (M < params->usedPairs) {
        c->set_error("ransac_umeyama: fewer matches than usedPairs (the reference's getRandomMatches would never return)");
        return RGBDFE_E_ARG;
    }
    // computeRANSACIteration (Ransac.cpp:449-454) for every reachable inlier ratio count / M, with the C library
    auto ransac_iters = [](double inlierRatio) {
        const double q = log(1 - 0.98) / log(1 - pow(inlierRatio, 3));
        if (!(q > -2147483649.0 && q < 2147483648.0))
            return (int)0x80000000; // what cvttsd2si yields for NaN / out-of-range, i.e. the reference binary's value
        return int(q);
    };
    const int iters0 = ransac_iters(0.20); // the ctor ignores params.iterationCount (Ransac.cpp:26)
    const int iters_min = ransac_iters(params->minimalInlierRatioThreshold);
    std::vector<int> tab((size_t)M + 1, 0);
    for (int k = 1; k <= M; k++)
        tab[k] = ransac_iters((double)(float(k) / float(M)));
    cudaSetDevice(c->device);
    int rc = ensure_pairs(c, 1);
    if (rc != RGBDFE_OK)
        return rc;
    rgbdfe_frame *Q = c->scratch, *T = c->scratch2;
    cudaStream_t s = c->stream;
    int* d_tab = nullptr;
    UmeState* d_state = nullptr;
    rgbdfe_umeyama_result* d_res = nullptr;
    CUDA_TRY(c, dmalloc(&d_tab, (size_t)M + 1));
    CUDA_TRY(c, dmalloc(&d_state, 1));
    CUDA_TRY(c, dmalloc(&d_res, 1));
    cudaError_t e = cudaMemcpyAsync(Q->xyz, p1, (size_t)n1 * 12, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(T->xyz, p2, (size_t)n2 * 12, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(c->d_matches, matches, (size_t)nm * sizeof(rgbdfe_dmatch), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_tab, tab.data(), ((size_t)M + 1) * sizeof(int), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        rc = rgbdfe::launch_ransac_umeyama(c, nm, Q->xyz, T->xyz, params->usedPairs, iters0, d_tab, iters_min, params->inlierThresholdEuclidean,
            params->errorVersion == 4, params->minimalInlierRatioThreshold, d_state, d_res);
        if (rc == RGBDFE_OK)
            e = cudaStreamSynchronize(s);
    }
    if (e == cudaSuccess && rc == RGBDFE_OK)
        e = cudaMemcpy(result, d_res, sizeof(*result), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && rc == RGBDFE_OK && inliers && result->n_inliers > 0) {
        std::vector<rgbdfe_dmatch> kept(M);
        std::vector<int> pos(result->n_inliers);
        e = cudaMemcpy(kept.data(), c->d_sorted, (size_t)M * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost);
        if (e == cudaSuccess)
            e = cudaMemcpy(pos.data(), c->d_inl_pos, pos.size() * sizeof(int), cudaMemcpyDeviceToHost);
        for (int i = 0; e == cudaSuccess && i < result->n_inliers && i < cap; i++)
            inliers[i] = kept[pos[i]];
        if (result->n_inliers > cap)
            rc = RGBDFE_E_CAPACITY;
    }
    cudaFree(d_tab), cudaFree(d_state), cudaFree(d_res);
    if (e != cudaSuccess) {
        c->set_error(std::string("ransac_umeyama: ") + cudaGetErrorString(e));
        return RGBDFE_E_CUDA;
    }
    return rc;
}

int rgbdfe_track(rgbdfe_ctx* c, rgbdfe_frame* ref, rgbdfe_frame* cur, float ratio, int discard_outliers, int iterations, unsigned min_inlier_th,
    float max_mahal, unsigned sample_size, int update_f2, int result_slot)
{
    if (!c || !ref || !cur || result_slot < 0 || result_slot >= c->result_cap)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    int rc = ensure_pairs(c, 1);
    if (rc != RGBDFE_OK)
        return rc;
    // Matcher::match(ref, cur): query = ref, train = cur (Tracking.cpp:127)
    MatchPair p = { ref->desc, cur->desc, ref->n, cur->n, ref->outlier, ref->xyz, cur->xyz, cur->outlier };
    if ((rc = set_pair(c, 0, p)) != RGBDFE_OK)
        return rc;
    if ((rc = launch_match(c, c->d_pairs, 1, c->kp_cap, c->kp_cap, ratio, discard_outliers, nullptr)) != RGBDFE_OK)
        return rc;
    Se3Params prm = make_params(c, iterations, min_inlier_th, max_mahal, sample_size);
    c->last_track_slot = result_slot, c->last_track_pairs = 1;
    return launch_ransac_se3(c, c->d_pairs, 0, 1, prm, update_f2, 0, 0, 0, 0, c->d_results + result_slot, nullptr);
}

int rgbdfe_track_sequence(rgbdfe_ctx* c, rgbdfe_frame* const* frames, int count, float ratio, int discard_outliers, int iterations,
    unsigned min_inlier_th, float max_mahal, unsigned sample_size, int update_f2, int first_result_slot)
{
    if (!c || !frames || count < 1 || count > CHAIN_MAX_FRAMES || first_result_slot < 0 || first_result_slot + count - 1 > c->result_cap)
        return RGBDFE_E_ARG;
    const int np = count - 1;
    if (np == 0)
        return RGBDFE_OK;
    cudaSetDevice(c->device);
    int rc = ensure_pairs(c, np);
    if (rc != RGBDFE_OK)
        return rc;
    FrameChain fc;
    for (int i = 0; i < count; i++) {
        if (!frames[i])
            return RGBDFE_E_ARG;
        fc.f[i] = *frames[i];
    }
    // A context bound to the latency partition sends the chain's throughput half (pair table + kNN, a grid
    // that fills every SM) to its stream on the throughput partition, ordered both ways with events.
    cudaStream_t chain_stream = c->stream;
    if (c->bulk_stream) {
        if (!c->bulk_evt)
            CUDA_TRY(c, cudaEventCreateWithFlags(&c->bulk_evt, cudaEventDisableTiming));
        CUDA_TRY(c, cudaEventRecord(c->bulk_evt, chain_stream));
        CUDA_TRY(c, cudaStreamWaitEvent(c->bulk_stream, c->bulk_evt, 0));
        c->stream = c->bulk_stream;
    }
    // pair p = Matcher::match(frames[p] /*ref, query*/, frames[p+1] /*cur, train*/) (Tracking.cpp:127)
    k_build_chain_pairs<<<1, CHAIN_MAX_FRAMES, 0, c->stream>>>(fc, np, c->d_pairs);
    c->launches++;
    // K7 for every pair of the chain at once: descriptors do not depend on the previous pair's outcome,
    // only the filter (outlier flags) and the RANSAC (rand() stream) do
    const int splits = rgbdfe::knn_splits_for(c, np);
    rc = rgbdfe::launch_knn(c, c->d_pairs, 0, np, c->kp_cap, splits);
    if (c->bulk_stream) {
        c->stream = chain_stream;
        if (rc == RGBDFE_OK) {
            CUDA_TRY(c, cudaEventRecord(c->bulk_evt, c->bulk_stream));
            CUDA_TRY(c, cudaStreamWaitEvent(chain_stream, c->bulk_evt, 0));
        }
    }
    if (rc != RGBDFE_OK)
        return rc;
    Se3Params prm = make_params(c, iterations, min_inlier_th, max_mahal, sample_size);
    rgbdfe::FusedFilter ff = { splits, ratio, discard_outliers };
    c->last_track_slot = first_result_slot, c->last_track_pairs = np;
    // The sequential part is two small kernels per pair, 128 dependent launches for a 64-frame batch, every one with the same
    // arguments batch after batch (the pair table is rebuilt above, workspaces and result slots are fixed).  Launches are a
    // host-side resource -- the driver serialises them per device context, about 3.5 us each, which with 32 sequences in flight
    // was a third of the whole step -- so the launch sequence of a chain is captured once per (length, parameters) and replayed
    // as ONE graph launch.  The first call with a given key runs un-captured (it also raises the kernels' shared-memory limits).
    const bool graphable = c->chain_graphs && np >= 4 && !c->profiling && !c->dbg_on;
    ChainGraphKey key;
    memset(&key, 0, sizeof(key)); // compared bytewise: padding must be defined
    key = ChainGraphKey { np, first_result_slot, splits, discard_outliers, iterations, (int)min_inlier_th, (int)sample_size, update_f2, c->ransac_first_chunk + 1000 * c->ransac_small_ctas,
        ratio, max_mahal, (void*)c->d_pairs };
    ChainGraph* cg = nullptr;
    if (graphable) {
        for (auto& g : c->chain_graph_cache)
            if (!memcmp(&g.key, &key, sizeof(key))) {
                cg = &g;
                break;
            }
        if (!cg) {
            if (c->chain_graph_cache.size() >= 8) { // a handful of chain shapes per context; drop the oldest
                if (c->chain_graph_cache.front().exec)
                    cudaGraphExecDestroy(c->chain_graph_cache.front().exec);
                c->chain_graph_cache.erase(c->chain_graph_cache.begin());
            }
            ChainGraph fresh;
            memset(&fresh.key, 0, sizeof(fresh.key));
            fresh.key = key;
            c->chain_graph_cache.push_back(fresh);
            cg = &c->chain_graph_cache.back();
        }
    }
    if (cg && cg->exec) {
        CUDA_TRY(c, cudaGraphLaunch(cg->exec, c->stream));
        c->launches += 2 * np; // the graph's kernel nodes
        return RGBDFE_OK;
    }
    const bool capture = cg && cg->seen > 0; // second call with this key
    if (capture)
        CUDA_TRY(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    for (int p = 0; p < np && rc == RGBDFE_OK; p++)
        rc = launch_ransac_se3(c, c->d_pairs, p, 1, prm, update_f2, 0, 0, 0, 0, c->d_results + first_result_slot + p, &ff);
    if (capture) {
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
        if (rc != RGBDFE_OK || e != cudaSuccess || !graph) {
            if (graph)
                cudaGraphDestroy(graph);
            cudaGetLastError();
            if (rc == RGBDFE_OK) {
                c->set_error(std::string("track_sequence: stream capture failed: ") + cudaGetErrorString(e));
                rc = RGBDFE_E_CUDA;
            }
            return rc;
        }
        const cudaError_t ei = cudaGraphInstantiate(&cg->exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) {
            cg->exec = nullptr;
            c->set_error(std::string("track_sequence: cudaGraphInstantiate: ") + cudaGetErrorString(ei));
            return RGBDFE_E_CUDA;
        }
        CUDA_TRY(c, cudaGraphLaunch(cg->exec, c->stream)); // nothing ran during the capture
    }
    if (cg)
        cg->seen++;
    return rc;
}

int rgbdfe_track_results(rgbdfe_ctx* c, int first, int count, rgbdfe_se3_result* results)
{
    if (!c || first < 0 || count < 0 || first + count > c->result_cap || (count > 0 && !results))
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    if (count > 0)
        CUDA_TRY(c, cudaMemcpyAsync(c->h_results + first, c->d_results + first, (size_t)count * sizeof(rgbdfe_se3_result), cudaMemcpyDeviceToHost,
                        c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    if (count > 0)
        memcpy(results, c->h_results + first, (size_t)count * sizeof(rgbdfe_se3_result));
    return RGBDFE_OK;
}

} // extern "C"

namespace {
// mvInliers of `count` pairs packed back to back: out[offsets[p] .. offsets[p + 1]) = sorted[pos[i]], offsets from the results
__global__ void k_pack_inliers(const rgbdfe_se3_result* __restrict__ res, const rgbdfe_dmatch* __restrict__ sorted_all, const int* __restrict__ pos_all,
    int match_cap, int pair0, rgbdfe_dmatch* __restrict__ out, int* __restrict__ offsets, int count)
{
    const int p = blockIdx.x;
    int off = 0;
    for (int i = 0; i < p; i++)
        off += res[i].n_inliers;
    const int n = res[p].n_inliers;
    const rgbdfe_dmatch* sorted = sorted_all + (size_t)(pair0 + p) * match_cap;
    const int* pos = pos_all + (size_t)(pair0 + p) * match_cap;
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        out[off + i] = sorted[pos[i]];
    if (threadIdx.x == 0) {
        offsets[p] = off;
        if (p == count - 1)
            offsets[count] = off + n;
    }
}
}

extern "C" int rgbdfe_track_results_inliers(rgbdfe_ctx* c, int first, int count, rgbdfe_se3_result* results, rgbdfe_dmatch* inliers, int inlier_cap,
    int* offsets)
{
    if (!c || first < 0 || count < 1 || count > CHAIN_MAX_FRAMES || first + count > c->result_cap || !results || !inliers || !offsets || inlier_cap < 0)
        return RGBDFE_E_ARG;
    const int pair0 = first - c->last_track_slot;
    if (pair0 < 0 || pair0 + count > c->last_track_pairs) {
        c->set_error("track_results_inliers: the slots must belong to the last rgbdfe_track / rgbdfe_track_sequence call of this context");
        return RGBDFE_E_STATE;
    }
    cudaSetDevice(c->device);
    const size_t total_cap = (size_t)count * c->kp_cap;
    if (total_cap > c->packed_cap) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_packed), cudaFreeHost(c->h_packed), cudaFree(c->d_dl);
        c->d_packed = nullptr, c->h_packed = nullptr, c->packed_cap = 0;
        CUDA_TRY(c, dmalloc(&c->d_packed, total_cap + 32)); // + room for the offsets behind the records
        CUDA_TRY(c, cudaMallocHost((void**)&c->h_packed, (total_cap + 32) * sizeof(rgbdfe_dmatch)));
        c->packed_cap = total_cap;
    }
    cudaStream_t s = c->stream;
    int* d_off = reinterpret_cast<int*>(c->d_packed + c->packed_cap);
    int* h_off = reinterpret_cast<int*>(c->h_packed + c->packed_cap);
    k_pack_inliers<<<count, 256, 0, s>>>(c->d_results + first, c->d_sorted, c->d_inl_pos, c->kp_cap, pair0, c->d_packed, d_off, count);
    c->launches++;
    CUDA_TRY(c, cudaMemcpyAsync(c->h_results + first, c->d_results + first, (size_t)count * sizeof(rgbdfe_se3_result), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaMemcpyAsync(h_off, d_off, (size_t)(count + 1) * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    memcpy(results, c->h_results + first, (size_t)count * sizeof(rgbdfe_se3_result));
    memcpy(offsets, h_off, (size_t)(count + 1) * sizeof(int));
    const int total = h_off[count];
    if (total > inlier_cap) {
        c->set_error("track_results_inliers: " + std::to_string(total) + " inliers exceed inlier_cap");
        return RGBDFE_E_CAPACITY;
    }
    if (total > 0) { // the exact number of records, known only now
        CUDA_TRY(c, cudaMemcpyAsync(c->h_packed, c->d_packed, (size_t)total * sizeof(rgbdfe_dmatch), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(c, cudaStreamSynchronize(s));
        memcpy(inliers, c->h_packed, (size_t)total * sizeof(rgbdfe_dmatch));
    }
    return RGBDFE_OK;
}

namespace rgbdfe {
cudaError_t raise_dyn_smem(int device, const void* kernel, size_t bytes)
{
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, size_t> have;
    std::lock_guard<std::mutex> lk(mu);
    size_t& cur = have[std::make_pair(device, kernel)];
    if (bytes <= cur)
        return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess)
        cur = bytes;
    return e;
}

void release_stream(rgbdfe_ctx* c)
{
    if (c->stream_owner) { // a sharer leaves
        auto& v = c->stream_owner->stream_sharers;
        v.erase(std::remove(v.begin(), v.end(), c), v.end());
        c->stream_owner = nullptr;
    } else if (c->owns_stream && !c->stream_sharers.empty()) { // an owner leaves while others still enqueue on the stream
        rgbdfe_ctx* heir = c->stream_sharers.front();
        heir->owns_stream = true;
        heir->stream_owner = nullptr;
        for (size_t i = 1; i < c->stream_sharers.size(); i++) {
            c->stream_sharers[i]->stream_owner = heir;
            heir->stream_sharers.push_back(c->stream_sharers[i]);
        }
        c->stream_sharers.clear();
    } else if (c->owns_stream && c->stream)
        cudaStreamDestroy(c->stream);
    c->stream = nullptr;
    c->owns_stream = false;
}
} // namespace rgbdfe

namespace rgbdfe { // helpers of this file used by verify.cu
int api_ensure_pairs(rgbdfe_ctx* c, int npairs) { return ensure_pairs(c, npairs); }
Se3Params api_make_params(rgbdfe_ctx* c, int iterations, unsigned min_inl, float max_mahal, unsigned sample) { return make_params(c, iterations, min_inl, max_mahal, sample); }
int api_pointer_kind(const void* p, const void** dev) { return pointer_kind(p, dev); }
}

extern "C" {

/* Developer aid (not part of the reference-facing surface): %globaltimer stamps of the chain kernels'
 * phases for pair slot 0 (see DBG_STAMP in ransac.cu).  on: 1/0; out: DBG_SLOTS values. */
int rgbdfe_debug_timeline(rgbdfe_ctx* c, int on, long long* out, int n)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    if (!c->d_dbg) {
        CUDA_TRY(c, dmalloc(&c->d_dbg, (size_t)DBG_SLOTS));
        CUDA_TRY(c, cudaMemset(c->d_dbg, 0, DBG_SLOTS * sizeof(long long)));
    }
    if (out && n > 0)
        CUDA_TRY(c, cudaMemcpy(out, c->d_dbg, (size_t)std::min(n, DBG_SLOTS) * sizeof(long long), cudaMemcpyDeviceToHost));
    CUDA_TRY(c, cudaMemset(c->d_dbg, 0, DBG_SLOTS * sizeof(long long)));
    c->dbg_on = on;
    return RGBDFE_OK;
}

int rgbdfe_profile_enable(rgbdfe_ctx* c, int on)
{
    if (!c)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    for (auto& e : c->prof_events) {
        c->prof_pool.push_back(e.a);
        c->prof_pool.push_back(e.b);
    }
    c->prof_events.clear();
    c->profiling = on != 0;
    return RGBDFE_OK;
}

int rgbdfe_profile_read(rgbdfe_ctx* c, double* ms_per_kernel, long long* launches_per_kernel, int n)
{
    if (!c || !ms_per_kernel || !launches_per_kernel || n < KID_COUNT)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < n; i++)
        ms_per_kernel[i] = 0, launches_per_kernel[i] = 0;
    for (auto& e : c->prof_events) {
        float ms = 0;
        CUDA_TRY(c, cudaEventElapsedTime(&ms, e.a, e.b));
        ms_per_kernel[e.id] += ms;
        launches_per_kernel[e.id]++;
    }
    return RGBDFE_OK;
}

const char* rgbdfe_profile_name(int id)
{
    static const char* names[KID_COUNT] = { "k_pyramid_level(x levels)", "k_fast_cells", "k_gauss7", "k_quadtree", "k_orient_describe", "k_knn2_umma",
        "k_ratio_filter", "k_se3_prepare", "k_se3_hypotheses", "k_se3_replay" };
    return (id >= 0 && id < KID_COUNT) ? names[id] : "";
}

} // extern "C"
