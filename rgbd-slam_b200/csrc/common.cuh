// Shared declarations of the rgbdfe CUDA library (sm_100a only).
#pragma once
#include "../../include/rgbdfe.h"

#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <string>
#include <vector>

#define RGBDFE_MAX_LEVELS 12
#define EDGE_THRESHOLD 19
#define CELL_CAP 256     // max NMS survivors kept per ~30 px FAST cell
#define CELL_MAX_DIM 72  // max cell ROI side (wCell + 6)
#define MAX_ITERS 512    // max RANSAC iterations per pair
#define QT_THREADS 256

// One pyramid level.  Offsets are relative to a frame slot's base pointer.
// Vertical resize program of PYR_ROWS consecutive padded rows of a level: the distinct source rows they read (byte
// offsets into the padded source level; unused entries repeat the last one) and, per output row, which two of them it
// blends (sel: 4 + 4 bits per row) with which coefficients (beta0 | beta1 << 16).
#define PYR_ROWS 4
#define PYR_SRC_MAX 8
struct PyrGroup {
    int off[PYR_SRC_MAX];
    unsigned beta[PYR_ROWS];
    unsigned sel;
    int nsrc, pad0, pad1;
};
struct LevelDesc {
    int w, h;            // un-padded size
    int pstride;         // padded row stride in bytes (>= w + 38, multiple of 64)
    int bstride;         // blurred row stride (multiple of 64)
    unsigned pyr_off;    // byte offset of the padded buffer (row -19, col -19)
    unsigned blur_off;
    int nCols, nRows, wCell, hCell; // FAST cell grid (ORBextractor.cpp:628-634)
    int maxBorderX, maxBorderY;     // w-16, h-16
    int cell_off;        // index of this level's first cell in the per-frame cell arrays
    int cand_off;        // first candidate slot of this level
    int cand_cap;
    int quota;           // mnFeaturesPerLevel
    int sel_off, sel_cap;
    int tab_x_off, tab_y_off; // first entry of the level in the padded column table / the row-group table (level > 0)
    float scale;         // mvScaleFactor[level]
    int patch_size;      // int(31 * scale)
};

struct OrbConfig {
    int nlevels;
    int nfeatures, iniTh, minTh;
    int w, h;            // image size the tables were built for
    int total_cells, total_cand, total_sel, max_quota;
    unsigned pyr_bytes, blur_bytes; // per frame slot
    int umax[16];
    int blur_tile_start[RGBDFE_MAX_LEVELS + 1]; // first k_gauss7 tile of each level; [nlevels] = total
    LevelDesc lv[RGBDFE_MAX_LEVELS];
};

// glibc TYPE_3 generator state, device-resident.
struct RngState {
    int32_t r[34];
    int32_t idx; // linear index of the next output
};

struct DepthLatch {
    int latched;
    int pad;
    double cov;
};

struct Se3Params {
    int iterations;
    unsigned min_inlier_th;
    float max_mahal;
    unsigned sample_size;
    double rcx, rcy; // raster covariances (SolverSE3.cpp:218-225), host tan()
};

struct SelKp { // keypoint selected by the quadtree, level coordinates
    short x, y;
    int score;
    float angle;
};

struct rgbdfe_frame {
    int cap;
    rgbdfe_keypoint* kps; // device
    uint8_t* desc;
    rgbdfe_point3f* xyz;
    uint8_t* outlier;
    int* n; // device count
    float2* pt_un; // mvKeysUn[i].pt (Core/Frame.cpp:251-281); equals kps[i].pt without distortion
    int* cell;     // Frame::posInGrid: x * FRAME_GRID_ROWS + y of the 64 x 48 grid, -1 outside (Core/Frame.cpp:75-89,231-241)
};

// Frame preprocessing parameters of one batched extraction, by value (N1: Core/Frame.cpp:47-48,251-316).
struct FramePrep {
    int undistort;            // distCoef[0] != 0 (Core/Frame.cpp:256)
    double fx, fy, cx, cy;    // the float intrinsics widened, as cv::undistortPoints sees the CV_32F matrix
    double ifx, ify;
    double k[5];              // k1 k2 p1 p2 k3
    float cxf, cyf, invfx, invfy; // uprojectCamera (float arithmetic)
    float minX, minY, gridWInv, gridHInv; // computeImageBounds / mfGridElement*Inv
    float maxX, maxY;
    float depth16_factor;     // > 0: the depth images are raw 16-bit, z = float(d) * factor
};

// Frames of one batched call, passed to kernels by value (<= 4 KB of kernel parameters).
struct FrameBatch {
    rgbdfe_frame f[64];
};

// Input images of one batched call (gray for K1, depth for K6), by value.  They point at the caller's buffers
// when those are device memory (no staging copy), else at the context's upload slots.
struct ImgPtrs {
    const void* p[64];
};

// Frames of one rgbdfe_track_sequence call, passed by value.
#define CHAIN_MAX_FRAMES 65
struct FrameChain {
    rgbdfe_frame f[CHAIN_MAX_FRAMES];
};

struct rgbdfe_kfdb {
    int n_kf, n_feat;
    uint8_t* desc;
    rgbdfe_point3f* xyz;
    int device;
};

// packed FAST candidate: x (12 bits) | y (12 bits) << 12 | score << 24
// cv::BORDER_REFLECT_101 index; |p| < 2*len always holds here (border 19 / kernel radius 3 vs len >= 20)
__host__ __device__ inline int reflect101(int p, int len)
{
    if (p < 0)
        p = -p;
    if (p >= len)
        p = 2 * (len - 1) - p;
    if (p < 0)
        p = -p;
    return p;
}
__host__ __device__ inline unsigned pack_cand(int x, int y, int s) { return (unsigned)x | ((unsigned)y << 12) | ((unsigned)s << 24); }
__host__ __device__ inline int cand_x(unsigned c) { return c & 0xfff; }
__host__ __device__ inline int cand_y(unsigned c) { return (c >> 12) & 0xfff; }
__host__ __device__ inline int cand_s(unsigned c) { return c >> 24; }

// One (query frame, train frame) pair handed to the match / RANSAC kernels.
struct MatchPair {
    const uint8_t* q_desc;
    const uint8_t* t_desc;
    const int* nq; // device counts
    const int* nt;
    const uint8_t* q_outlier; // may be null
    const rgbdfe_point3f* q_xyz;
    const rgbdfe_point3f* t_xyz;
    uint8_t* t_outlier;       // F2 flags updated when update_f2 (may be null)
};

// Result of one RANSAC hypothesis (one executed iteration of SolverSE3.cpp:54-103).
struct HypOut {
    int count;  // vRefinedMatches.size()
    int pad;
    double err; // refinedError
    float T[16];
};

// Sequential accept / skip / break state of the RANSAC loop, replayed on the device.
struct ReplayState {
    int n;          // the reference's loop variable (with its += 10 skips)
    int k;          // executed iterations so far (= hypotheses consumed)
    int done;
    int validIters;
    int bestHyp;    // -1 none
    int bestCount;
    float rmse;
    int run;        // RANSAC loop runs for this pair
    int nm;         // matches seen
    int rng_calls;  // rand() calls consumed by the executed iterations
    int sampled_k;     // iterations whose samples have been drawn so far
    int sampled_calls; // rand() calls those samples consumed
};

// State handed from k_ume_prepare to the later K12 kernels.
struct UmeState {
    int M;           // matches after the depth gate (Ransac.cpp:72-83)
    int total_calls; // rand() calls made while drawing all initial iterations
};

// A captured chain (rgbdfe_track_sequence): the 2 x np kernel launches of the sequential part as one graph.
struct ChainGraphKey {
    int np, first_slot, splits, discard, iterations, min_inl, sample, update_f2, first_chunk;
    float ratio, max_mahal;
    void* pairs; // the workspace generation
};
struct ChainGraph {
    ChainGraphKey key;
    cudaGraphExec_t exec = nullptr;
    int seen = 0;
};

struct rgbdfe_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t bulk_stream = nullptr; // RGBDFE_PART_LATENCY: stream on the throughput partition for the chain's kNN (partition.cu)
    cudaEvent_t order_evt = nullptr; // rgbdfe_stream_wait
    cudaEvent_t bulk_evt = nullptr;  // stream <-> bulk_stream ordering
    int img_channels = 1;            // channels the image upload slots are sized for (3 after the first BGR batch)
    int depth_upload_of_8 = 0;       // option: of every 8 extraction calls this many upload a pinned depth map although host_zero_copy is on
    unsigned extract_calls = 0;
    bool host_zero_copy = true;      // a pinned host depth map is read in place by K6 (keypoint sectors only) instead of being uploaded
    std::string err;
    long long launches = 0;
    int max_w = 0, max_h = 0, max_feats = 0, max_batch = 1;
    int kp_cap = 0; // per-frame keypoint capacity (max_feats + slack)

    // ORB configuration
    bool configured = false;
    int nfeatures = 1000, nlevels = 8, iniTh = 20, minTh = 7;
    float scaleFactorF = 1.2f;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    OrbConfig cfg;          // host copy (valid for cfg.w x cfg.h)
    OrbConfig* d_cfg = nullptr;

    // extraction workspaces, max_batch frame slots each
    uint8_t* d_img = nullptr;   size_t img_slot = 0;
    float* d_depth = nullptr;   size_t depth_slot = 0;
    uint8_t* d_pyr = nullptr;   size_t pyr_slot = 0;
    uint8_t* d_blur = nullptr;  size_t blur_slot = 0;
    int* d_cell_count = nullptr; unsigned* d_cell_list = nullptr; int cell_slot = 0;
    unsigned* d_cand = nullptr; int* d_label = nullptr; int cand_slot = 0;
    int* d_cand_count = nullptr; // [slot][level]
    SelKp* d_sel = nullptr; int sel_slot = 0;
    int* d_sel_count = nullptr;  // [slot][level]
    // resize tables of levels > 0 over PADDED coordinates (ring and pitch padding resolved on the host):
    // x: (source column offset in the padded source row, alpha0 | alpha1 << 16) per padded column of the row pitch;
    // y: one PyrGroup per PYR_ROWS padded rows
    int2* d_tab_xpad = nullptr; PyrGroup* d_tab_group = nullptr;
    int pyr_nsrc = 0; // most distinct source rows any row group of the current geometry reads
    int tab_x_cap = 0, tab_y_cap = 0;
    uint8_t* h_img = nullptr;   // pinned staging
    float* h_depth = nullptr;
    rgbdfe_frame* scratch = nullptr; // frame used by the host-buffer extract API
    rgbdfe_frame* scratch2 = nullptr;
    int last_extract_slot = 0;

    // match / RANSAC workspaces for pair_cap pairs
    int pair_cap = 0;
    int knn_splits = 8;
    // option "overlap_blur": K3 (blur, needs only the pyramid) runs on a side stream beside K4 (quadtree, latency-bound)
    bool owns_stream = true; // false after rgbdfe_share_stream
    rgbdfe_ctx* stream_owner = nullptr;   // whose stream this context enqueues on (rgbdfe_share_stream)
    std::vector<rgbdfe_ctx*> stream_sharers; // contexts enqueuing on THIS context's stream
    int overlap_blur = 0;
    cudaStream_t side_stream = nullptr;
    cudaEvent_t side_fork = nullptr, side_join = nullptr;
    int knn_waves = 6; // train-range splits are chosen to give about this many waves of kNN CTAs (option "knn_waves")
    bool knn_umma_ready = false; // k_knn2_umma's dynamic shared memory attribute set on this context's device
    int knn_umma = 1; // K7 on tcgen05 (match_umma.cu): 1 = k_knn2_umma (default, the faster one as measured), 2 = the warp-specialised persistent variant, 0 = mma.sync (option "knn_umma")
    int knn_dbg = 0;  // developer aid (option "knn_dbg"): bit 0 skips the fold, 1 the train expansion, 2 the multiplications -- wrong results, stage timing only
    int sm_count = 0; // SMs of the device (grid of the persistent kNN kernel)
    MatchPair* d_pairs = nullptr;   MatchPair* h_pairs = nullptr; // pinned staging
    int4* d_partial = nullptr;      // [pairs][splits][kp_cap]
    rgbdfe_dmatch* d_matches = nullptr; // [pairs][kp_cap]  Matcher::match output (query order)
    rgbdfe_dmatch* d_sorted = nullptr;  // [pairs][kp_cap]  std::sort'ed copy used by RANSAC
    rgbdfe_point3f* d_P1 = nullptr; rgbdfe_point3f* d_P2 = nullptr; float* d_w = nullptr; // matched points / TFC weights, sorted order
    int* d_nmatch = nullptr;        // [pairs]
    int* d_samples = nullptr;       // [pairs][MAX_ITERS][8]
    int* d_rng_used = nullptr;      // [pairs][MAX_ITERS] rand() calls consumed after iteration k
    HypOut* d_hyp = nullptr;        // [pairs][MAX_ITERS]
    unsigned* d_hyp_mask = nullptr; // [pairs][MAX_ITERS][mask_words]
    int mask_words = 0;
    ReplayState* d_replay = nullptr; // [pairs]
    int* d_hyp_ready = nullptr;     // [pairs][MAX_ITERS] hypothesis published (reset by k_se3_prepare)
    int* d_pair_lock = nullptr;     // [pairs] replay lock
    rgbdfe_se3_result* d_results = nullptr; int result_cap = 0;
    rgbdfe_se3_result* h_results = nullptr; // pinned
    int* d_inl_pos = nullptr;       // [pairs][kp_cap] positions (in the sorted match list) of the winning inliers
    int* d_const_n = nullptr;       // small device ints for host-API counts
    RngState* d_rng = nullptr;      // context stream
    RngState* d_pair_rng = nullptr; // [pairs] per-pair streams (batched)
    DepthLatch* d_latch = nullptr;
    double rcx = 0, rcy = 0;
    int ransac_verify_cap = 1;      // many-pairs path: 1 = register-capped small hypothesis CTAs, 0 = 512-thread CTAs, -n = 512-thread CTAs below n pairs per launch
    int ransac_small_ctas = 0;      // option: tracking chains use the register-capped 128 / 256-thread hypothesis CTAs of the verification path
    int ransac_first_chunk = 32;    // hypotheses of the first launch of a latency-mode RANSAC (rgbdfe_set_option)
    // K2 state (orb_image.cu): host copy of the kernel's parameter block (tensor maps of the pyramid levels) and the cell table
    void* fast_grid = nullptr; void* d_fast_cells = nullptr; void* d_fast_maps = nullptr; bool fast_ready = false;
    const void* fast_key_pyr = nullptr; int fast_key_w = 0, fast_key_h = 0, fast_key_levels = 0, fast_key_slots = 0;
    struct CvOrbWorkspace* cvorb = nullptr; // the cv::ORB branch of Extractor (row N3, orb_cv.cu), allocated by rgbdfe_cvorb_configure
    int extractor = 0;                      // RGBDFE_EXTRACTOR_ORB2 / RGBDFE_EXTRACTOR_ORB: which one the extraction entry points run
    struct ProjWorkspace* proj = nullptr; // Matcher::projectionMatch buffers, allocated on first use (projection.cu)
    bool chain_graphs = true; // option "chain_graphs": replay a chain's launch sequence as one CUDA graph
    std::vector<ChainGraph> chain_graph_cache;
    int last_track_slot = 0, last_track_pairs = 0; // result slots of the last rgbdfe_track / rgbdfe_track_sequence (pair p -> slot + p)
    rgbdfe_dmatch* d_packed = nullptr; rgbdfe_dmatch* h_packed = nullptr; size_t packed_cap = 0; // rgbdfe_track_results_inliers staging
    unsigned char* d_dl = nullptr; size_t dl_bytes = 0; // rgbdfe_frames_download_async staging
    struct VerifyState* verify = nullptr; // sharded verification: comm stream, result slots, query staging (verify.cu)
    bool latch_known = false;       // the host has seen the depth-covariance latch set (batched verification requires it)
    long long* d_dbg = nullptr;     // %globaltimer stamps written by the chain kernels when debug timing is on (rgbdfe_debug_timeline)
    int dbg_on = 0;

    // optional per-kernel timing (CUDA events on `stream`), see rgbdfe_profile_*
    bool profiling = false;
    struct ProfEvt { int id; cudaEvent_t a, b; };
    std::vector<ProfEvt> prof_events;
    std::vector<cudaEvent_t> prof_pool;

    void set_error(const std::string& s) { err = s; }
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute and setting it REPLACES the value: the largest
// size requested so far is kept per (device, kernel) for the whole process, under a lock (contexts of several devices
// and host threads share the kernels), and the attribute only ever grows.  The caller has made c->device current.
namespace rgbdfe {
cudaError_t raise_dyn_smem(int device, const void* kernel, size_t bytes); // api.cu
}
template <class K>
inline cudaError_t ensure_dyn_smem(rgbdfe_ctx* c, K kernel, size_t bytes)
{
    return rgbdfe::raise_dyn_smem(c->device, (const void*)kernel, bytes);
}

#define CUDA_TRY(ctx, expr)                                                       \
    do {                                                                          \
        cudaError_t _e = (expr);                                                  \
        if (_e != cudaSuccess) {                                                  \
            (ctx)->set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            return RGBDFE_E_CUDA;                                                 \
        }                                                                         \
    } while (0)

#define DBG_SLOTS 1024
#ifdef __CUDACC__
__device__ __forceinline__ long long dbg_now()
{
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define DBG_STAMP(ptr, slot)                 \
    do {                                     \
        if ((ptr) && (slot) < DBG_SLOTS)     \
            (ptr)[(slot)] = dbg_now();       \
    } while (0)
#endif

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

// kernel ids for rgbdfe_profile_read
enum { KID_PYRAMID = 0, KID_FAST, KID_BLUR, KID_QUADTREE, KID_DESCRIBE, KID_KNN, KID_FILTER, KID_SE3_PREPARE, KID_SE3_HYP, KID_SE3_REPLAY, KID_COUNT };

// RAII helper: times one kernel (or launch group) when ctx->profiling is on.
struct ProfScope {
    rgbdfe_ctx* c;
    cudaEvent_t a = nullptr, b = nullptr;
    int id;
    static cudaEvent_t get(rgbdfe_ctx* c)
    {
        if (!c->prof_pool.empty()) {
            cudaEvent_t e = c->prof_pool.back();
            c->prof_pool.pop_back();
            return e;
        }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
    ProfScope(rgbdfe_ctx* ctx, int kid) : c(ctx), id(kid)
    {
        if (c->profiling) {
            a = get(c);
            b = get(c);
            cudaEventRecord(a, c->stream);
        }
    }
    ~ProfScope()
    {
        if (a) {
            cudaEventRecord(b, c->stream);
            c->prof_events.push_back({ id, a, b });
        }
    }
};

namespace rgbdfe {
// orb_image.cu
int launch_pyramid(rgbdfe_ctx* c, int count, const ImgPtrs& gray, int gray_stride, int channels);
int launch_blur(rgbdfe_ctx* c, int count);
int launch_fast(rgbdfe_ctx* c, int count);
void free_fast_state(rgbdfe_ctx* c);
// orb_keypoints.cu
int launch_quadtree(rgbdfe_ctx* c, int count);
int launch_describe(rgbdfe_ctx* c, int count, const FrameBatch& frames, bool unproject, const ImgPtrs& depth, int depth_stride, const FramePrep& prep,
    bool depth_on_host);
// match.cu
int launch_match(rgbdfe_ctx* c, const MatchPair* d_pairs, int npairs, int max_nq, int max_nt, float ratio, int discard, int* d_knn_out);
int launch_knn(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits);
int knn_splits_for(const rgbdfe_ctx* c, int npairs);
int launch_knn_umma(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits);
// projection.cu
void free_proj_workspace(rgbdfe_ctx* c);
// orb_cv.cu
void free_cvorb_workspace(rgbdfe_ctx* c);
int cvorb_configure(rgbdfe_ctx* c, int nfeatures, float scaleFactor, int nlevels);
int launch_cvorb(rgbdfe_ctx* c, int count, int w, int h, const ImgPtrs& gray, int gray_stride, const FrameBatch& frames, bool unproject,
    const ImgPtrs& depth, int depth_stride, const FramePrep& prep);
int cvorb_stage_level(rgbdfe_ctx* c, int level, int which, uint8_t* dst, int dst_stride, int* w, int* h);
int cvorb_stage_harris(rgbdfe_ctx* c, int level, rgbdfe_keypoint* out, int cap, int* n);
// ransac.cu
// Pairs occupy workspace slots [slot0, slot0 + npairs); results go to d_results[0 .. npairs).  With `ff` the
// Matcher::match filter (K8) runs inside k_se3_prepare on the kNN partials of the same slots.
struct FusedFilter {
    int splits;
    float ratio;
    int discard_outliers;
    // per-pair streams of a batch of query groups: pair p seeds srand(seed + first_pair_index + (p / seed_group) * seed_stride + p % seed_group);
    // 0 = one group (seed + first_pair_index + p)
    int seed_group = 0, seed_stride = 0;
};
int launch_ransac_se3(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, const Se3Params& prm, int update_f2, int per_pair_rng,
    unsigned seed, int first_pair_index, int min_matches, rgbdfe_se3_result* d_results, const FusedFilter* ff);
// api.cu
// The context stops using c->stream (it is about to get another one, or to be destroyed): a sharer leaves its owner; an
// owner with sharers hands the stream to the first of them (it stays alive as long as anyone enqueues on it), an owner
// without sharers destroys it.
void release_stream(rgbdfe_ctx* c);
int api_ensure_pairs(rgbdfe_ctx* c, int npairs);
Se3Params api_make_params(rgbdfe_ctx* c, int iterations, unsigned min_inl, float max_mahal, unsigned sample);
int api_pointer_kind(const void* p, const void** dev); // 0 pageable host, 1 device, 2 pinned host
// verify.cu
void free_verify_state(rgbdfe_ctx* c);
// ransac_umeyama.cu
int launch_ransac_umeyama(rgbdfe_ctx* c, int nm, const rgbdfe_point3f* d_xyz1, const rgbdfe_point3f* d_xyz2, int used_pairs, int iters0,
    const int* d_iters_tab, int iters_min, double th0, int adaptive, double min_ratio, UmeState* d_state, rgbdfe_umeyama_result* d_res);
}
