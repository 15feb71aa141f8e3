/*
 * rgbdfe.h — C ABI of the B200-native RGB-D tracking front end.
 *
 * Drop-in boundary for the frame-to-frame tracking hot path of
 * toniortiz/rgbd-slam: ORB extraction -> Hamming kNN + ratio test -> 3D-3D SE3
 * RANSAC, plus batched loop-pair verification.  The reference has no FFI; its
 * boundary is four C++ classes.  The host shims under rgbd-slam_b200/host/
 * re-create those classes on top of exactly these entry points (INTEGRATION.md
 * shows the binding).  Every function returns 0 on success and a negative
 * RGBDFE_E_* code on error; rgbdfe_last_error() returns the message.  No CPU
 * fallback exists: without a CUDA device rgbdfe_create() fails.
 *
 * Plain pointers and sizes only; the caller owns every host buffer, the
 * context owns all device memory, streams, the RNG stream and the
 * depth-covariance latch.
 *
 * Layout-compatible PODs: rgbdfe_keypoint == cv::KeyPoint (28 B),
 * rgbdfe_dmatch == cv::DMatch (16 B), rgbdfe_point3f == cv::Point3f (12 B).
 */
#ifndef RGBDFE_H
#define RGBDFE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RGBDFE_OK 0
#define RGBDFE_E_ARG (-1)      /* bad argument */
#define RGBDFE_E_CAPACITY (-2) /* output / context capacity too small */
#define RGBDFE_E_CUDA (-3)     /* CUDA runtime error, see rgbdfe_last_error */
#define RGBDFE_E_STATE (-4)    /* call order / configuration error */

typedef struct rgbdfe_ctx rgbdfe_ctx;
typedef struct rgbdfe_frame rgbdfe_frame; /* device-resident Frame data (Core/Frame.h:141-155,195) */
typedef struct rgbdfe_kfdb rgbdfe_kfdb;   /* device-resident keyframe shard for batched verification */

typedef struct {
    float x, y, size, angle, response;
    int octave, class_id;
} rgbdfe_keypoint;

typedef struct {
    int queryIdx, trainIdx, imgIdx;
    float distance;
} rgbdfe_dmatch;

typedef struct {
    float x, y, z;
} rgbdfe_point3f;

/* Public results of RansacSE3 (Solver/SolverSE3.h:54-56) + bookkeeping. */
typedef struct {
    int ok;             /* return value of RansacSE3::compute */
    float rmse;         /* RansacSE3::rmse */
    float T21[16];      /* RansacSE3::mT21, row-major 4x4 */
    int n_inliers;      /* mvInliers.size() */
    int iters_executed; /* RANSAC iterations the sequential reference would have run */
    int n_matches;      /* size of the match list the solver saw */
} rgbdfe_se3_result;

/* Ransac::parameters subset that is executable in the reference (Solver/Ransac.h:21-29). */
typedef struct {
    int errorVersion; /* 0 EUCLIDEAN_ERROR, 4 ADAPTIVE_ERROR */
    double inlierThresholdEuclidean;
    double minimalInlierRatioThreshold;
    int minimalNumberOfMatches;
    int usedPairs;
} rgbdfe_umeyama_params;

typedef struct {
    int ok; /* return value of Ransac::compute (= !T.isIdentity()) */
    float T[16];
    int n_inliers;
    int iters_executed;
} rgbdfe_umeyama_result;

/* ---- context ---------------------------------------------------------- */
/* max_w/max_h: largest gray image; max_feats: nfeatures upper bound;
 * max_batch: frames processed per batched call (>= 1). */
int rgbdfe_create(rgbdfe_ctx** ctx, int device, int max_w, int max_h, int max_feats, int max_batch);
void rgbdfe_destroy(rgbdfe_ctx* ctx);
const char* rgbdfe_last_error(const rgbdfe_ctx* ctx);
/* cudaStream_t the context launches on (for callers that time with events). */
void* rgbdfe_stream(rgbdfe_ctx* ctx);
/* number of kernels launched by the context so far (bench gpu_launches). */
long long rgbdfe_launch_count(const rgbdfe_ctx* ctx);
int rgbdfe_sync(rgbdfe_ctx* ctx);
/* Tuning knobs that never change results.  "ransac_first_chunk" (default 32): how many RANSAC iterations the
 * first hypothesis launch of a single pair evaluates speculatively; the rest follow in a second launch only
 * if the reference's loop has not ended by then.  Small values suit many concurrent sequences on one GPU.
 * "ransac_small_ctas" (default 0): the hypothesis CTAs of tracking chains are 128 / 256 threads with registers capped at 128
 * (the verification path's shape) instead of 512 threads, which at 128 registers take an SM's whole register file: slower
 * for one chain, but an SM that hosts one still has room for the extraction kernels of other sequences (+4 % frames/s
 * with 32 sequences on one GPU).  "ransac_verify_cap" (default 1; 0 = the uncapped shapes, for A/B runs) is the same choice
 * for the many-pairs path of rgbdfe_verify_*.
 * "host_zero_copy" (default 1): see rgbdfe_frames_extract.  "depth_upload_of_8" (default 0): of every 8 extraction calls
 * of the context, that many upload a pinned depth map whole although host_zero_copy is on (in-place reads are bound by the
 * PCIe read rate, whole uploads by the PCIe bandwidth; a mix uses both); "depth_upload_phase": where in that cycle of 8
 * the context starts, so that several contexts take turns.
 * "knn_umma" (default 1): the Hamming kNN of Matcher::match on tcgen05: 1 = k_knn2_umma (two CTAs per SM, every warp does
 * every job), 2 = the warp-specialised persistent variant k_knn2_umma_ws (query tile in tensor memory; measured 10 % slower, see
 * DESIGN.md), 0 = the mma.sync kernel (k_knn2_hamming); 0 and 2 are kept for A/B measurements — all three are held to the
 * same parity tests.
 * "knn_waves" (default 6): the train range of a pair is split over CTAs so that a kNN launch has about this many
 * waves of CTAs (latency of single pairs against redundant work for whole chains). */
int rgbdfe_set_option(rgbdfe_ctx* ctx, const char* name, int value);
/* Pipelining across contexts of one device (e.g. one context extracts frame batch n+1 while another tracks
 * batch n; rgbdfe_frame handles are plain device buffers and may be used by either): everything enqueued on
 * `waiter` after this call runs after everything enqueued on `signaler` before it.  No host synchronisation. */
int rgbdfe_stream_wait(rgbdfe_ctx* waiter, rgbdfe_ctx* signaler);
/* high != 0: the context's stream gets the device's highest priority (for the latency-bound match/RANSAC
 * chain next to a throughput-bound extraction context).  Synchronises the context. */
int rgbdfe_set_stream_priority(rgbdfe_ctx* ctx, int high);
/* `ctx` gives up its own stream and enqueues on `owner`'s from now on (same device).  The stream lives as long as any
 * context enqueues on it: when its owner is destroyed or takes another stream first, ownership passes to a sharer.  For many throughput-bound extraction
 * contexts on one GPU: their big grids gain nothing from a stream each, while every extra stream aliases onto the 32
 * hardware work queues and can serialise an unrelated latency-bound chain behind them.  Synchronises `ctx`. */
int rgbdfe_share_stream(rgbdfe_ctx* ctx, rgbdfe_ctx* owner);
/* SM partitioning (CUDA green contexts) for the same split, stronger than priorities: the device's SMs are
 * divided once per process into a latency partition of about `latency_sms` SMs (the driver rounds to its
 * granularity, 8 SMs on sm_90+) and a throughput partition with the rest; the sizes actually provisioned are
 * returned.  A second call for the same device returns the existing split.  RGBDFE_E_STATE: the driver has
 * no green contexts (callers fall back to rgbdfe_set_stream_priority). */
int rgbdfe_device_partition(int device, int latency_sms, int* latency_got, int* throughput_got);
/* Moves the context's stream into a partition of its device.  RGBDFE_PART_THROUGHPUT: extraction contexts.
 * RGBDFE_PART_LATENCY: tracking contexts — the filter/sort/RANSAC chain of rgbdfe_track_sequence runs on the
 * latency partition at the highest stream priority while its kNN (one grid over every pair of the chain) is
 * sent to the throughput partition.  RGBDFE_PART_NONE: back to an ordinary stream.  Synchronises the context;
 * rgbdfe_stream() changes.  Results never depend on the binding. */
#define RGBDFE_PART_NONE 0
#define RGBDFE_PART_THROUGHPUT 1
#define RGBDFE_PART_LATENCY 2
int rgbdfe_bind_partition(rgbdfe_ctx* ctx, int role);
/* Per-kernel timing with CUDA events on the context stream (bench.py roofline).
 * enable(1) starts collecting, read() synchronises and returns the summed
 * milliseconds and launch-group counts per kernel id (n >= 10 entries);
 * profile_name(id) names the kernel.  Off by default: no events are recorded. */
int rgbdfe_profile_enable(rgbdfe_ctx* ctx, int on);
int rgbdfe_profile_read(rgbdfe_ctx* ctx, double* ms_per_kernel, long long* launches_per_kernel, int n);
const char* rgbdfe_profile_name(int id);

/* Developer aid: device-side phase time stamps (ns) of the match/RANSAC chain kernels for the first pair
 * of the last launch; on = 1 starts recording, out (may be NULL) receives up to n of 512 slots. */
int rgbdfe_debug_timeline(rgbdfe_ctx* ctx, int on, long long* out, int n);

/* ---- Extractor / ORBextractor ---------------------------------------- */
/* Extractor::setParameters (Features/Extractor.cpp:24-48) -> ORBextractor ctor
 * (Features/ORBextractor.cpp:348-406). */
int rgbdfe_orb_configure(rgbdfe_ctx* ctx, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
/* Extractor::getScaleFactors / getInverseScaleFactors / getScaleSigmaSquares /
 * getInverseScaleSigmaSquares (Features/Extractor.cpp:111-158) + per-level quotas.
 * Each array has nlevels entries; any pointer may be NULL. */
int rgbdfe_orb_tables(rgbdfe_ctx* ctx, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* quota);
/* ORBextractor::operator() (Features/ORBextractor.cpp:706-766): gray u8 image in,
 * keypoints + 32-byte descriptors out.  *n receives the keypoint count. */
int rgbdfe_orb_extract(rgbdfe_ctx* ctx, const uint8_t* gray, int w, int h, int stride, rgbdfe_keypoint* kps, uint8_t* desc, int cap, int* n);
/* ORBextractor::mvImagePyramid[level] of the last extract, including the 19 px
 * border ring (Features/ORBextractor.cpp:773-797).  dst is (w+38) x (h+38). */
int rgbdfe_orb_pyramid_level(rgbdfe_ctx* ctx, int level, uint8_t* dst, int dst_stride, int* w, int* h);
/* Stage outputs of the last extract, for stage-by-stage parity tests:
 * per-cell FAST candidates before the quadtree (coordinates relative to the
 * detection window, as at ORBextractor.cpp:666-668) and the blurred level. */
int rgbdfe_orb_stage_candidates(rgbdfe_ctx* ctx, int level, rgbdfe_keypoint* out, int cap, int* n);
int rgbdfe_orb_stage_blurred(rgbdfe_ctx* ctx, int level, uint8_t* dst, int dst_stride);

/* ---- Extractor(ORB, ORB): the cv::ORB branch --------- */
/* Extractor::eType values the library implements: ORB2 = the in-tree ORBextractor (rgbdfe_orb_configure),
 * ORB = OpenCV's cv::ORB with Harris scoring (Features/Extractor.cpp:163-166,216-218). */
#define RGBDFE_EXTRACTOR_ORB2 0
#define RGBDFE_EXTRACTOR_ORB 1
/* Extractor(ORB, ORB).setParameters(nfeatures, scaleFactor, nlevels, ...) (Features/Extractor.cpp:24-48,163-166):
 * from now on every extraction entry point of this context (rgbdfe_orb_extract, rgbdfe_frames_extract,
 * rgbdfe_frames_extract_raw) runs Extractor::detectAndCompute's non-ORB2 branch (Features/Extractor.cpp:55-59):
 * cv::ORB::create(nfeatures, scaleFactor, nlevels)->detect, retainBest(nfeatures), cv::ORB::create()->compute.
 * rgbdfe_orb_configure switches back.  scaleFactor must be 1.2f: the reference's descriptor object is a default
 * cv::ORB, whose pyramid only coincides with the detector's at that value.  Gray input only. */
int rgbdfe_cvorb_configure(rgbdfe_ctx* ctx, int nfeatures, float scaleFactor, int nlevels);
/* Stage outputs of slot 0 of the last cv::ORB extraction, for stage-by-stage parity tests: the level image
 * (which = 0), the blurred level (1) or the non-max-suppressed FAST score map (2), tight w x h; and the keypoints
 * that entered the second retainBest (after retainBest(2 N_l) and HarrisResponses), level coordinates. */
int rgbdfe_cvorb_stage_level(rgbdfe_ctx* ctx, int level, int which, uint8_t* dst, int dst_stride, int* w, int* h);
int rgbdfe_cvorb_stage_harris(rgbdfe_ctx* ctx, int level, rgbdfe_keypoint* out, int cap, int* n);

/* ---- Matcher ---------------------------------------------------------- */
/* Matcher::match (Features/Matcher.cpp:106-139): kNN k=2 Hamming + ratio test +
 * unique-train + outlier / valid-depth filters.  q_outlier may be NULL. */
int rgbdfe_match_knn2(rgbdfe_ctx* ctx, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt, const uint8_t* q_outlier,
    const rgbdfe_point3f* q_xyz, const rgbdfe_point3f* t_xyz, float ratio, int discard_outliers, rgbdfe_dmatch* out, int cap, int* n);

/* Matcher::match for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same result as
 * rgbdfe_match_knn2 (the mma.sync kNN kernel + the ratio / unique-train / outlier / depth filter), but every pointer
 * is a DEVICE pointer supplied by the caller, the call is asynchronous on `stream` (a cudaStream_t; NULL = the
 * context's stream) and it calls no cudaMalloc / cudaFree / allocating routine: the kernels' scratch (pair
 * descriptor, counts, top-2 partials) is one static __device__ array of 256 KB + 256 B that every GPU loading the
 * library holds until exit.  Limit: nq, nt <= 4096, beyond that RGBDFE_E_CAPACITY with a message in
 * rgbdfe_last_error.  out_dev receives at most cap matches in query order, *n_dev = min(number found, cap).
 * All calls of a process share that array per device: calls that overlap on the GPU (two streams without an
 * ordering between them) race; calls on one stream, or ordered by events, are safe.  ctx supplies the device,
 * the default stream and the error string only. */
int rgbdfe_match_knn2_noalloc(rgbdfe_ctx* ctx, void* stream, const uint8_t* q_desc_dev, int nq, const uint8_t* t_desc_dev, int nt,
    const uint8_t* q_outlier_dev, const rgbdfe_point3f* q_xyz_dev, const rgbdfe_point3f* t_xyz_dev, float ratio, int discard_outliers,
    rgbdfe_dmatch* out_dev, int cap, int* n_dev);

/* ---- Random (System/Random.cpp) --------------------------------------- */
/* The context owns one glibc-rand()-compatible stream (TYPE_3).  state is 35
 * int32: 34 ring words + the linear output index. */
int rgbdfe_rng_seed(rgbdfe_ctx* ctx, unsigned seed);
/* One rand() draw from the context stream (what Random::randomInt / randomValue consume,
 * System/Random.h:13-18, System/Random.cpp:15-20). */
int rgbdfe_rng_next(rgbdfe_ctx* ctx, int* value);
int rgbdfe_rng_get_state(rgbdfe_ctx* ctx, int32_t* state35);
int rgbdfe_rng_set_state(rgbdfe_ctx* ctx, const int32_t* state35);

/* ---- RansacSE3 -------------------------------------------------------- */
/* depthCovariance latch (Solver/SolverSE3.cpp:282-287): function-local statics
 * latched from the first depth seen in the process. */
int rgbdfe_depth_cov_get(rgbdfe_ctx* ctx, int* latched, double* cov);
int rgbdfe_depth_cov_set(rgbdfe_ctx* ctx, int latched, double cov);
/* RansacSE3::compute (Solver/SolverSE3.cpp:23-133) on host arrays.  p1/p2 are
 * Frame::mvKeys3Dc of F1/F2.  inliers (may be NULL) receives mvInliers. The
 * Frame side effects (setOutlier / setInlier / setPose) stay with the shim. */
int rgbdfe_ransac_se3(rgbdfe_ctx* ctx, const rgbdfe_point3f* p1, int n1, const rgbdfe_point3f* p2, int n2, const rgbdfe_dmatch* matches,
    int nm, int iterations, unsigned min_inlier_th, float max_mahal, unsigned sample_size, rgbdfe_se3_result* result,
    rgbdfe_dmatch* inliers, int cap);

/* ---- Ransac (Umeyama) ------------------------------------------------- */
/* Ransac::estimateTransformation (Solver/Ransac.cpp:61-181). */
int rgbdfe_ransac_umeyama(rgbdfe_ctx* ctx, const rgbdfe_point3f* p1, int n1, const rgbdfe_point3f* p2, int n2,
    const rgbdfe_dmatch* matches, int nm, const rgbdfe_umeyama_params* params, rgbdfe_umeyama_result* result,
    rgbdfe_dmatch* inliers, int cap);

/* ---- device-resident frames (tracking pipeline) ----------------------- */
int rgbdfe_frame_create(rgbdfe_ctx* ctx, rgbdfe_frame** frame);
void rgbdfe_frame_destroy(rgbdfe_ctx* ctx, rgbdfe_frame* frame);
/* Frame ctor hot part (Core/Frame.cpp:51,91-117): extract ORB from gray, then
 * back-project every keypoint with the metric depth map (float32, 0 = invalid)
 * and pinhole intrinsics; outlier flags reset to false.  Asynchronous on the
 * context stream; the buffers must stay valid and unchanged until rgbdfe_sync().
 * gray/depth may be device pointers (read in place, no staging copy) or host pointers
 * (ideally pinned; uploaded into the context's slots, images that follow each other in
 * memory as one transfer).  A pinned host DEPTH image is not uploaded: the back-projection
 * reads only the sectors under the keypoints over PCIe, in place (as the reference reads
 * mImDepth only at the keypoints); rgbdfe_set_option "host_zero_copy" 0 uploads it whole.  All images of one call must be
 * of the same kind.
 * count frames are processed as one batch (count <= max_batch). */
int rgbdfe_frames_extract(rgbdfe_ctx* ctx, rgbdfe_frame* const* frames, int count, const uint8_t* const* gray, int w, int h, int stride,
    const float* const* depth, int depth_stride_elems, float fx, float fy, float cx, float cy);
/* The same with the RAW depth image of the dataset (16-bit, as cv::imread(..., IMREAD_UNCHANGED) returns it):
 * Frame's `imDepth.convertTo(mImDepth, CV_32F, mDepthMapFactor)` (Core/Frame.cpp:48, mDepthMapFactor = 1.0f / DepthMapFactor,
 * Core/RGBDcamera.cpp:21) is applied on the device where the keypoints are back-projected — z = float(d16) * depth_map_factor,
 * OpenCV's float convert-scale — so a frame costs 2 instead of 4 bytes of depth per pixel on the way to the GPU. */
int rgbdfe_frames_extract_raw(rgbdfe_ctx* ctx, rgbdfe_frame* const* frames, int count, const uint8_t* const* gray, int w, int h, int stride,
    const uint16_t* const* depth16, int depth_stride_elems, float depth_map_factor, float fx, float fy, float cx, float cy);
/* N1 -- the whole of Frame::Frame's image work (Core/Frame.cpp:31-73) for count frames: cvtColor(BGR2GRAY) when
 * channels == 3 (OpenCV's 8-bit fixed point; channels == 1 takes a gray image), depth convertTo(CV_32F, factor),
 * extractFeatures, undistortKeyPoints (cv::undistortPoints(p, K, distCoef, noArray(), K): five iterations in double;
 * skipped when dist[0] == 0 as at Frame.cpp:256), assignFeaturesToGrid (the 64 x 48 cell of every undistorted
 * keypoint; bounds from the undistorted image corners, Frame.cpp:283-316) and uprojectCamera (depth at the raw
 * position, back-projection of the undistorted one). */
typedef struct rgbdfe_camera {
    float fx, fy, cx, cy;
    float dist[5];          /* k1 k2 p1 p2 k3 (RGBDcamera::distCoef) */
    float depth_map_factor; /* mDepthMapFactor = 1.0f / DepthMapFactor (Core/RGBDcamera.cpp:21) */
} rgbdfe_camera;
int rgbdfe_frames_from_images(rgbdfe_ctx* ctx, rgbdfe_frame* const* frames, int count, const uint8_t* const* color, int channels, int w, int h,
    int stride, const uint16_t* const* depth16, int depth_stride_elems, const rgbdfe_camera* cam);
/* mvKeysUn[i].pt (2 floats per keypoint) and the grid cell x * 48 + y (-1: outside the grid) of a frame; every
 * extraction entry point fills them (without distortion mvKeysUn == mvKeys).  Pointers may be NULL. */
int rgbdfe_frame_download_undistorted(rgbdfe_ctx* ctx, rgbdfe_frame* frame, float* pts_un, int* grid_cell, int cap, int* n);
/* ---- Matcher::projectionMatch (Features/Matcher.cpp:35-104), scope row N2 ---- */
/* Everything the function reads besides the keypoint arrays: F1's camera (:37-40) and world pose (mRwc, mOw of
 * Frame::unprojectWorld, Core/Frame.cpp:317-327), F2's pose (getRotation / getTranslation, :42-43), the static image
 * bounds and grid cell sizes of Frame (mnMinX.., mfGridElementWidthInv/HeightInv, Core/Frame.cpp:63-66,283-316)
 * and the search radius th (Matcher.h:26, default 5).  Matrices are row-major CV_32F as in the reference. */
typedef struct rgbdfe_projection {
    float fx, fy, cx, cy;
    float Rwc1[9], Ow1[3];
    float Rcw2[9], tcw2[3];
    float min_x, max_x, min_y, max_y;
    float grid_w_inv, grid_h_inv;
    float th;
} rgbdfe_projection;
/* Frame::computeImageBounds + the grid cell sizes for a camera and image size: bounds = {mnMinX, mnMaxX, mnMinY,
 * mnMaxY}, grid_inv = {mfGridElementWidthInv, mfGridElementHeightInv}.  Host arithmetic only. */
int rgbdfe_camera_bounds(const rgbdfe_camera* cam, int w, int h, float bounds[4], float grid_inv[2]);
/* projectionMatch on host arrays (what the Matcher shim binds): q_* = F1 (mDescriptors, mvKeys3Dc, mvbOutlier as
 * bytes or NULL), t_* = F2 (mDescriptors, mvKeys3Dc, mvKeysUn[i].pt as 2 floats per keypoint; the mGrid cell of each
 * is recomputed on the device with Frame::posInGrid).  out receives DMatch(i1, bestIdx2, bestDist) in query order,
 * imgIdx -1 as the cv::DMatch constructor leaves it; *n is the number found (RGBDFE_E_CAPACITY when > cap). */
int rgbdfe_projection_match(rgbdfe_ctx* ctx, const uint8_t* q_desc, int nq, const rgbdfe_point3f* q_xyz, const uint8_t* q_outlier,
    const uint8_t* t_desc, int nt, const rgbdfe_point3f* t_xyz, const float* t_pts_un, const rgbdfe_projection* proj, rgbdfe_dmatch* out,
    int cap, int* n);
/* The same for npairs (F1, F2) pairs of device-resident frames in one pass (descriptors, 3-D points, outlier flags,
 * undistorted points and grid cells are the ones the extraction left on the device).  out is [npairs][cap_per_pair],
 * n is [npairs].  Blocking. */
int rgbdfe_frames_projection_match(rgbdfe_ctx* ctx, rgbdfe_frame* const* f1, rgbdfe_frame* const* f2, const rgbdfe_projection* proj, int npairs,
    rgbdfe_dmatch* out, int cap_per_pair, int* n);
/* Diagnostic: rounds the ordered-claim fixpoint of the last projection call took per pair (1 + the longest chain of
 * keypoints that lost their first choice to an earlier one). */
int rgbdfe_projection_rounds(rgbdfe_ctx* ctx, int* rounds, int npairs);
/* Resident host-side frames for the class shims: Matcher::match(ref, cur) and RansacSE3::compute(F1, F2) are called on
 * the same two Frames one after the other (Tracking.cpp:125-130) and `cur` becomes `ref` of the next call, so a caller
 * that keeps a device copy per Frame uploads every descriptor / 3-D point set once instead of three times.
 * frame_upload: descriptors (n x 32 B) + mvKeys3Dc (n x 12 B) into a frame made by rgbdfe_frame_create; asynchronous.
 * frames_match = rgbdfe_match_knn2 and frames_ransac_se3 = rgbdfe_ransac_se3 on such frames (nq / nt / n1 / n2 are the
 * counts uploaded); q_outlier (may be NULL) are F1's CURRENT mvbOutlier flags, uploaded on every call because the
 * host changes them (Frame::setOutlier / setInlier). */
int rgbdfe_frame_upload(rgbdfe_ctx* ctx, rgbdfe_frame* frame, const uint8_t* desc, const rgbdfe_point3f* xyz, int n);
int rgbdfe_frames_match(rgbdfe_ctx* ctx, rgbdfe_frame* ref, rgbdfe_frame* cur, int nq, int nt, const uint8_t* q_outlier, float ratio,
    int discard_outliers, rgbdfe_dmatch* out, int cap, int* n);
int rgbdfe_frames_ransac_se3(rgbdfe_ctx* ctx, rgbdfe_frame* f1, rgbdfe_frame* f2, int n1, int n2, const rgbdfe_dmatch* matches, int nm, int iterations,
    unsigned min_inlier_th, float max_mahal, unsigned sample_size, rgbdfe_se3_result* result, rgbdfe_dmatch* inliers, int cap);
/* Copy a frame's data back (any pointer may be NULL). */
int rgbdfe_frame_download(rgbdfe_ctx* ctx, rgbdfe_frame* frame, rgbdfe_keypoint* kps, uint8_t* desc, rgbdfe_point3f* xyz,
    uint8_t* outlier, int cap, int* n);
/* Tracking::visualOdometry core (System/Tracking.cpp:125-130): Matcher(ratio)
 * .match(ref, cur, discardOutliers) then RansacSE3(...).compute(ref, cur,
 * matches, updateF2) entirely on the device; cur's outlier flags are updated as
 * SolverSE3.cpp:38-42,119-123 do.  Asynchronous: result_slot indexes a
 * context-owned result ring read back with rgbdfe_track_results(). */
int rgbdfe_track(rgbdfe_ctx* ctx, rgbdfe_frame* ref, rgbdfe_frame* cur, float ratio, int discard_outliers, int iterations,
    unsigned min_inlier_th, float max_mahal, unsigned sample_size, int update_f2, int result_slot);
/* The same for a tracked sequence: pair i = (frames[i], frames[i+1]), i = 0 .. count-2, each the
 * Matcher::match + RansacSE3::compute of rgbdfe_track, executed in sequence order (pair i sees the
 * outlier flags pair i-1 left on frames[i] and continues the context's rand() stream), results in
 * slots first_result_slot + i.  The Hamming kNN of all pairs runs as one launch ahead of the
 * sequential part.  count <= 65.  Asynchronous. */
int rgbdfe_track_sequence(rgbdfe_ctx* ctx, rgbdfe_frame* const* frames, int count, float ratio, int discard_outliers, int iterations,
    unsigned min_inlier_th, float max_mahal, unsigned sample_size, int update_f2, int first_result_slot);
/* Synchronise and copy results [first, first+count) of the ring to the host. */
int rgbdfe_track_results(rgbdfe_ctx* ctx, int first, int count, rgbdfe_se3_result* results);
/* The same plus RansacSE3::mvInliers (Solver/SolverSE3.h:55) of every pair: the slots must be those of the LAST
 * rgbdfe_track / rgbdfe_track_sequence call of this context (the inlier sets live in its pair workspace).  The
 * DMatch records of pair p are inliers[offsets[p] .. offsets[p + 1]) in the reference's order (offsets has count + 1
 * entries); RGBDFE_E_CAPACITY when the total exceeds inlier_cap.  count <= 65. */
int rgbdfe_track_results_inliers(rgbdfe_ctx* ctx, int first, int count, rgbdfe_se3_result* results, rgbdfe_dmatch* inliers, int inlier_cap,
    int* offsets);
/* What Frame keeps on the host after its constructor (mvKeys, mDescriptors, mvKeys3Dc) for count device-resident frames,
 * enqueued on the context stream without synchronising: row i of every output is frame i, `cap` entries wide (whole
 * rows travel, n[i] says how many are valid, the rest is unspecified); count <= 64; pinned host buffers make the copies
 * asynchronous.  Any of kps / desc /
 * xyz may be NULL.  The buffers are complete after rgbdfe_sync(). */
int rgbdfe_frames_download_async(rgbdfe_ctx* ctx, rgbdfe_frame* const* frames, int count, rgbdfe_keypoint* kps, uint8_t* desc, rgbdfe_point3f* xyz,
    int cap, int* n);

/* ---- batched loop-pair verification (PoseGraph.cpp:135-154,260-284) ---- */
/* A keyframe shard: n_kf keyframes x n_feat features each, resident on the
 * device.  desc: n_kf*n_feat*32 B, xyz: n_kf*n_feat*12 B (host or device pointers). */
int rgbdfe_kfdb_create(rgbdfe_ctx* ctx, rgbdfe_kfdb** db, int n_kf, int n_feat, const uint8_t* desc, const rgbdfe_point3f* xyz);
void rgbdfe_kfdb_destroy(rgbdfe_ctx* ctx, rgbdfe_kfdb* db);
/* For every keyframe k of the shard: Matcher(ratio).match(KF_k, query) then
 * RansacSE3(iterations, min_inlier_th, max_mahal, sample).compute(KF_k, query,
 * m, false) with the per-pair stream srand(seed + first_pair_index + k)
 * (SURVEY 8e).  Pairs with fewer than min_matches matches report ok = 0 without
 * running RANSAC (PoseGraph.cpp:142-144).  results: n_kf entries (host).  If
 * results_dev != NULL the same structs are also left at that device address
 * (for an NCCL gather without a host hop). */
int rgbdfe_verify_batch(rgbdfe_ctx* ctx, rgbdfe_kfdb* db, const uint8_t* q_desc, const rgbdfe_point3f* q_xyz, int nq, float ratio,
    int min_matches, int iterations, unsigned min_inlier_th, float max_mahal, unsigned sample_size, unsigned seed, int first_pair_index,
    rgbdfe_se3_result* results, void* results_dev);

/* ---- sharded verification over the GPUs of one box (BASELINE configs[3] and [4]) ------------------ */
/* The keyframe database is block-sharded in keyframe order: rank r of n_ranks owns keyframes [lo, hi) as
 * rgbdfe_shard_range returns them (block sizes differ by at most one) and keeps them resident as one rgbdfe_kfdb.
 * Every rank is handed the same query frame(s); each verifies its block exactly as rgbdfe_verify_batch does
 * (the loop body of PoseGraph.cpp:135-154,260-284) and ONE ncclAllGather of the 84-byte results per call rebuilds
 * the full list on every rank -- the only exchange step of the path.  Pair (query q, keyframe k) draws from the
 * stream srand(seed + q * n_kf_total + k), k the GLOBAL keyframe index, so results do not depend on n_ranks.
 * The first success in candidate order (PoseGraph.cpp:283) stays a host pick over the returned list.
 * The depth-covariance latch must be set (rgbdfe_depth_cov_set, or a tracked pair before): RGBDFE_E_STATE otherwise,
 * here and in rgbdfe_verify_batch -- the reference latches it from the first pair processed; a batch has no order. */
void rgbdfe_shard_range(int n_kf_total, int rank, int n_ranks, int* lo, int* hi);
typedef struct rgbdfe_verify_params {
    float ratio;            /* Matcher(ratio) */
    int min_matches;        /* PoseGraph.cpp:142-144 */
    int iterations;         /* RansacSE3(iterations, min_inlier_th, max_mahal, sample_size) */
    unsigned min_inlier_th;
    float max_mahal;
    unsigned sample_size;
    unsigned seed;
} rgbdfe_verify_params;
/* NCCL is resolved at run time (dlopen of libnccl.so.2, or the path in RGBDFE_NCCL_LIB), so the library also loads
 * where NCCL is absent.  unique_id: 128 bytes made by rank 0 and handed to every rank by the application (any
 * bootstrap: file, socket, MPI, torch.distributed); comm_create is ncclCommInitRank on the context's device and
 * returns an ncclComm_t.  An ncclComm_t the application created itself from the same NCCL library works as well. */
int rgbdfe_nccl_unique_id(void* id128);
int rgbdfe_nccl_comm_create(rgbdfe_ctx* ctx, int n_ranks, int rank, const void* id128, void** comm);
int rgbdfe_nccl_comm_destroy(void* comm);
/* Asynchronous half: verifies n_queries query frames (q_desc [n_queries][nq][32], q_xyz [n_queries][nq]; host
 * memory, or device memory that is then read in place) against this rank's block on the context stream, then --
 * on a communication stream of the context, so that the next enqueue's kernels overlap it -- the all-gather
 * (skipped for n_ranks == 1, nccl_comm may be NULL) and the copy of the gathered results to pinned host memory.
 * slot in [0, 4) names the result buffers used; a slot must be collected before it is enqueued again. */
int rgbdfe_verify_sharded_enqueue(rgbdfe_ctx* ctx, rgbdfe_kfdb* db, void* nccl_comm, int n_ranks, int rank, int n_kf_total, const uint8_t* q_desc,
    const rgbdfe_point3f* q_xyz, int nq, int n_queries, const rgbdfe_verify_params* params, int slot);
/* Waits for the slot and writes results[q * n_kf_total + k] for every query q and GLOBAL keyframe k (NULL: wait only). */
int rgbdfe_verify_sharded_collect(rgbdfe_ctx* ctx, int slot, rgbdfe_se3_result* results);
/* Blocking form: enqueue on slot 0 + collect. */
int rgbdfe_verify_sharded(rgbdfe_ctx* ctx, rgbdfe_kfdb* db, void* nccl_comm, int n_ranks, int rank, int n_kf_total, const uint8_t* q_desc,
    const rgbdfe_point3f* q_xyz, int nq, int n_queries, const rgbdfe_verify_params* params, rgbdfe_se3_result* results);
/* Host arithmetic only: the gathered layout [rank][query][ceil(n_kf_total / n_ranks)] -> [query][n_kf_total]. */
int rgbdfe_gathered_to_keyframe_order(const rgbdfe_se3_result* gathered, int n_ranks, int n_queries, int n_kf_total, rgbdfe_se3_result* out);

/* ---- LoopDetector (PlaceRecognition/LoopDetector.cpp:22-84), scope row N4: candidate keyframes for detectLoop ---- */
/* The bag-of-words vectors are INPUTS (DBoW3::BowVector = ascending (word id, double weight) pairs, made by
 * Frame::computeBoW -> DBoW3::Vocabulary::transform, Core/Frame.cpp:243-249; DBoW3 and its vocabulary file are not part
 * of the reference tree).  bowdb_add = LoopDetector::add (:22-26): keyframes keep their insertion order, which is the
 * order of every inverted-file list; *index (may be NULL) receives the keyframe's index in the database. */
typedef struct rgbdfe_bowdb rgbdfe_bowdb;
int rgbdfe_bowdb_create(rgbdfe_ctx* ctx, rgbdfe_bowdb** db, int max_kf, int max_words_per_kf);
void rgbdfe_bowdb_destroy(rgbdfe_ctx* ctx, rgbdfe_bowdb* db);
int rgbdfe_bowdb_add(rgbdfe_ctx* ctx, rgbdfe_bowdb* db, int kf_id, const int* words, const double* weights, int n, int* index);
/* LoopDetector::obtainCandidates(pKF) (:28-84) for the keyframe at query_index: `connected` = indices of
 * pKF->getConnectedKFs() (empty: no candidates, :33-35), interval = mnInterval.  Vocabulary::score is DBoW3's L1 scoring
 * (the ORB vocabulary's), accumulated in double over the common words in ascending order.  Returns at most five
 * database indices in the reference's order (the encounter order of the inverted-file walk, or std::sort by
 * Frame::CompareBoWscore when more than five qualify) with their scores.  The mnSharingWord marks persist in the
 * database between calls, as the Frame member does (a second call for the same keyframe finds nothing, :56-57). */
int rgbdfe_loop_candidates(rgbdfe_ctx* ctx, rgbdfe_bowdb* db, int query_index, const int* connected, int n_connected, int interval, int* candidates,
    double* scores, int cap, int* n);

#ifdef __cplusplus
}
#endif
#endif /* RGBDFE_H */
