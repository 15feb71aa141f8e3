// Batched and sharded loop-pair verification: PoseGraph::createLocalEdges / detectLoop (Solver/PoseGraph.cpp:128-155,
// 245-287) run one Matcher::match + RansacSE3::compute per candidate keyframe in a serial host loop.  Here a query
// frame is verified against every keyframe of a device-resident block in one pass (K7 for all pairs, then K8-K11 per
// pair), and the block may be one shard of a keyframe database spread over the GPUs of a box: every rank verifies
// its block and ONE ncclAllGather of the 84-byte per-pair results rebuilds the list (the only exchange step of the
// path).  NCCL is resolved at run time (dlopen of libnccl.so.2) so that the library loads on hosts without it.
//
// Data flow of one enqueue (context stream s, communication stream cs, result slot b):
//   s : [upload queries] -> per query { pair table (1 kernel) -> kNN -> prepare(+filter) -> hypotheses } -> event computed[b]
//   cs: wait computed[b] -> ncclAllGather(block -> all) -> D2H into pinned memory -> event done[b]
// A later enqueue runs its kernels on s while cs still gathers; slot b's block is rewritten only after done[b].
#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "common.cuh"

using namespace rgbdfe;

#define VERIFY_SLOTS 4
#define VERIFY_GROUP_PAIRS 4096 // pairs per pass when several query frames are verified in one call

namespace {

struct VerifySlot {
    rgbdfe_se3_result* d_block = nullptr; // [n_queries][block_max], this rank's results
    rgbdfe_se3_result* d_all = nullptr;   // [n_ranks][n_queries][block_max]
    rgbdfe_se3_result* h_all = nullptr;   // pinned copy of d_all (of d_block when n_ranks == 1)
    size_t cap_block = 0, cap_all = 0;    // records
    cudaEvent_t computed = nullptr, done = nullptr;
    bool pending = false, used = false;
    int n_ranks = 0, n_queries = 0, block_max = 0, n_kf_total = 0;
};

} // namespace

struct VerifyState {
    cudaStream_t comm_stream = nullptr;
    VerifySlot slot[VERIFY_SLOTS];
    uint8_t* d_qdesc = nullptr; // staging for host queries
    rgbdfe_point3f* d_qxyz = nullptr;
    size_t q_cap = 0; // descriptors
};

namespace {

template <class T>
cudaError_t dmalloc(T** p, size_t n)
{
    return cudaMalloc((void**)p, n * sizeof(T) > 0 ? n * sizeof(T) : 1);
}

// ---- NCCL, resolved at run time ---------------------------------------------------------------------------------
struct NcclId {
    char internal[128]; // ncclUniqueId
};
struct NcclApi {
    void* lib = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    std::string err;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

const NcclApi* nccl_api(std::string* why)
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib)
        return &g_nccl;
    if (!g_nccl.err.empty()) {
        *why = g_nccl.err;
        return nullptr;
    }
    const char* names[3] = { getenv("RGBDFE_NCCL_LIB"), "libnccl.so.2", "libnccl.so" };
    void* h = nullptr;
    for (const char* n : names)
        if (n && *n && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL)))
            break;
    if (!h) {
        g_nccl.err = std::string("NCCL not found (dlopen libnccl.so.2; set RGBDFE_NCCL_LIB): ") + (dlerror() ? dlerror() : "");
        *why = g_nccl.err;
        return nullptr;
    }
    g_nccl.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    g_nccl.GetUniqueId = (int (*)(NcclId*))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(h, "ncclCommInitRank");
    g_nccl.CommDestroy = (int (*)(void*))dlsym(h, "ncclCommDestroy");
    g_nccl.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(h, "ncclAllGather");
    if (!g_nccl.GetErrorString || !g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllGather) {
        g_nccl.err = "the NCCL library lacks a required symbol";
        *why = g_nccl.err;
        dlclose(h);
        return nullptr;
    }
    g_nccl.lib = h;
    return &g_nccl;
}

// Matcher(ratio).match(KF_k, cur): query side = keyframe k of the block, train side = the current (query) frame
// (PoseGraph.cpp:139-141).  Pair p of a group of query frames is (query p / block_max, keyframe p % block_max); the
// pairs past the end of a short block are empty (both counts 0).  Also writes the device counts the pairs point at.
__global__ void k_build_verify_pairs(MatchPair* __restrict__ dst, int np, int block_max, int n_block, const uint8_t* kf_desc,
    const rgbdfe_point3f* kf_xyz, int n_feat, const uint8_t* q_desc, const rgbdfe_point3f* q_xyz, int nq, int* counts)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p == 0)
        counts[0] = n_feat, counts[1] = nq, counts[2] = 0;
    if (p >= np)
        return;
    const int q = p / block_max, k = p % block_max;
    const bool real = k < n_block;
    const int kk = real ? k : 0;
    dst[p] = MatchPair { kf_desc + (size_t)kk * n_feat * 32, q_desc + (size_t)q * nq * 32, real ? counts : counts + 2, real ? counts + 1 : counts + 2, nullptr,
        kf_xyz + (size_t)kk * n_feat, q_xyz + (size_t)q * nq, nullptr };
}

// The depth-covariance latch (SolverSE3.cpp:282-287) must be set before a batch: the reference latches it from the
// first pair it processes, a batch has no first pair.
int require_latch(rgbdfe_ctx* c)
{
    if (c->latch_known)
        return RGBDFE_OK;
    DepthLatch l;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(&l, c->d_latch, sizeof(l), cudaMemcpyDeviceToHost));
    if (!l.latched) {
        c->set_error("batched verification needs the depth-covariance latch set (rgbdfe_depth_cov_set, or track a pair first): "
                     "the reference latches it from the first pair processed and a batch has no pair order");
        return RGBDFE_E_STATE;
    }
    c->latch_known = true;
    return RGBDFE_OK;
}

// n_q query frames against every keyframe of the block as ONE batch of n_q * block_max pairs; results
// d_res[q * block_max + k].  Pair (q, k) seeds srand(seed + q * seed_stride + first_pair_index + k).  Asynchronous.
int verify_core(rgbdfe_ctx* c, const rgbdfe_kfdb* db, const uint8_t* dq_desc, const rgbdfe_point3f* dq_xyz, int nq, int n_q, int block_max,
    const rgbdfe_verify_params& vp, unsigned seed, int first_pair_index, int seed_stride, rgbdfe_se3_result* d_res)
{
    const int np = n_q * block_max;
    k_build_verify_pairs<<<(np + 255) / 256, 256, 0, c->stream>>>(c->d_pairs, np, block_max, db->n_kf, db->desc, db->xyz, db->n_feat, dq_desc, dq_xyz, nq,
        c->d_const_n);
    c->launches++;
    const int splits = knn_splits_for(c, np);
    int rc = launch_knn(c, c->d_pairs, 0, np, db->n_feat, splits);
    if (rc != RGBDFE_OK)
        return rc;
    const Se3Params prm = api_make_params(c, vp.iterations, vp.min_inlier_th, vp.max_mahal, vp.sample_size);
    const FusedFilter ff = { splits, vp.ratio, 1, n_q > 1 ? block_max : 0, seed_stride }; // K8 runs inside k_se3_prepare
    return launch_ransac_se3(c, c->d_pairs, 0, np, prm, 0, 1, seed, first_pair_index, vp.min_matches, d_res, &ff);
}

int check_query_args(rgbdfe_ctx* c, const rgbdfe_kfdb* db, const void* q_desc, const void* q_xyz, int nq)
{
    if (!c || !db || !q_desc || !q_xyz || nq < 1)
        return RGBDFE_E_ARG;
    if (nq > c->kp_cap || db->n_feat > c->kp_cap) {
        c->set_error("verify: descriptor count exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    if (db->device != c->device) {
        c->set_error("verify: the keyframe block lives on another device than the context");
        return RGBDFE_E_ARG;
    }
    return RGBDFE_OK;
}

// Device-visible query arrays: used in place when they are device memory, else staged (stream-ordered upload).
int stage_queries(rgbdfe_ctx* c, VerifyState* vs, const uint8_t* q_desc, const rgbdfe_point3f* q_xyz, size_t n, const uint8_t** dq_desc,
    const rgbdfe_point3f** dq_xyz)
{
    const void *dd = nullptr, *dx = nullptr;
    if (api_pointer_kind(q_desc, &dd) == 1 && api_pointer_kind(q_xyz, &dx) == 1) {
        *dq_desc = (const uint8_t*)dd, *dq_xyz = (const rgbdfe_point3f*)dx;
        return RGBDFE_OK;
    }
    if (n > vs->q_cap) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        cudaFree(vs->d_qdesc), cudaFree(vs->d_qxyz);
        vs->d_qdesc = nullptr, vs->d_qxyz = nullptr, vs->q_cap = 0;
        CUDA_TRY(c, dmalloc(&vs->d_qdesc, n * 32));
        CUDA_TRY(c, dmalloc(&vs->d_qxyz, n));
        vs->q_cap = n;
    }
    CUDA_TRY(c, cudaMemcpyAsync(vs->d_qdesc, q_desc, n * 32, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(vs->d_qxyz, q_xyz, n * sizeof(rgbdfe_point3f), cudaMemcpyHostToDevice, c->stream));
    *dq_desc = vs->d_qdesc, *dq_xyz = vs->d_qxyz;
    return RGBDFE_OK;
}

int ensure_state(rgbdfe_ctx* c)
{
    if (c->verify)
        return RGBDFE_OK;
    VerifyState* vs = new (std::nothrow) VerifyState();
    if (!vs)
        return RGBDFE_E_ARG;
    c->verify = vs;
    int lo = 0, hi = 0;
    CUDA_TRY(c, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CUDA_TRY(c, cudaStreamCreateWithPriority(&vs->comm_stream, cudaStreamNonBlocking, hi));
    for (auto& sl : vs->slot) {
        CUDA_TRY(c, cudaEventCreateWithFlags(&sl.computed, cudaEventDisableTiming));
        CUDA_TRY(c, cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    return RGBDFE_OK;
}

int ensure_slot(rgbdfe_ctx* c, VerifySlot& sl, size_t block_records, size_t all_records)
{
    if (block_records > sl.cap_block || all_records > sl.cap_all) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->verify->comm_stream));
        cudaFree(sl.d_block), cudaFree(sl.d_all), cudaFreeHost(sl.h_all);
        sl.d_block = sl.d_all = sl.h_all = nullptr;
        sl.cap_block = sl.cap_all = 0;
        CUDA_TRY(c, dmalloc(&sl.d_block, block_records));
        CUDA_TRY(c, dmalloc(&sl.d_all, all_records));
        CUDA_TRY(c, cudaMallocHost((void**)&sl.h_all, all_records * sizeof(rgbdfe_se3_result)));
        sl.cap_block = block_records, sl.cap_all = all_records;
    }
    return RGBDFE_OK;
}

} // namespace

namespace rgbdfe {
void free_verify_state(rgbdfe_ctx* c)
{
    VerifyState* vs = c->verify;
    if (!vs)
        return;
    if (vs->comm_stream)
        cudaStreamSynchronize(vs->comm_stream);
    for (auto& sl : vs->slot) {
        cudaFree(sl.d_block), cudaFree(sl.d_all), cudaFreeHost(sl.h_all);
        if (sl.computed)
            cudaEventDestroy(sl.computed);
        if (sl.done)
            cudaEventDestroy(sl.done);
    }
    cudaFree(vs->d_qdesc), cudaFree(vs->d_qxyz);
    if (vs->comm_stream)
        cudaStreamDestroy(vs->comm_stream);
    delete vs;
    c->verify = nullptr;
}
} // namespace rgbdfe

extern "C" {

int rgbdfe_kfdb_create(rgbdfe_ctx* c, rgbdfe_kfdb** out, int n_kf, int n_feat, const uint8_t* desc, const rgbdfe_point3f* xyz)
{
    if (!c || !out || n_kf < 1 || n_feat < 1 || !desc || !xyz)
        return RGBDFE_E_ARG;
    *out = nullptr;
    if (n_feat > c->kp_cap) {
        c->set_error("kfdb: n_feat exceeds max_feats + 64");
        return RGBDFE_E_CAPACITY;
    }
    cudaSetDevice(c->device);
    rgbdfe_kfdb* db = new (std::nothrow) rgbdfe_kfdb();
    if (!db)
        return RGBDFE_E_ARG;
    db->n_kf = n_kf, db->n_feat = n_feat, db->desc = nullptr, db->xyz = nullptr, db->device = c->device;
    const size_t N = (size_t)n_kf * n_feat;
    cudaError_t e = dmalloc(&db->desc, N * 32);
    if (e == cudaSuccess)
        e = dmalloc(&db->xyz, N);
    if (e == cudaSuccess)
        e = cudaMemcpy(db->desc, desc, N * 32, cudaMemcpyDefault); // host or device source
    if (e == cudaSuccess)
        e = cudaMemcpy(db->xyz, xyz, N * sizeof(rgbdfe_point3f), cudaMemcpyDefault);
    if (e != cudaSuccess) {
        c->set_error(std::string("kfdb_create: ") + cudaGetErrorString(e));
        cudaFree(db->desc), cudaFree(db->xyz);
        delete db;
        return RGBDFE_E_CUDA;
    }
    *out = db;
    return RGBDFE_OK;
}

void rgbdfe_kfdb_destroy(rgbdfe_ctx* c, rgbdfe_kfdb* db)
{
    if (!db)
        return;
    if (c) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
    }
    cudaFree(db->desc);
    cudaFree(db->xyz);
    delete db;
}

int rgbdfe_verify_batch(rgbdfe_ctx* c, rgbdfe_kfdb* db, const uint8_t* q_desc, const rgbdfe_point3f* q_xyz, int nq, float ratio, int min_matches,
    int iterations, unsigned min_inlier_th, float max_mahal, unsigned sample_size, unsigned seed, int first_pair_index,
    rgbdfe_se3_result* results, void* results_dev)
{
    int rc = check_query_args(c, db, q_desc, q_xyz, nq);
    if (rc != RGBDFE_OK)
        return rc;
    if (!results && !results_dev)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    if ((rc = require_latch(c)) != RGBDFE_OK || (rc = ensure_state(c)) != RGBDFE_OK)
        return rc;
    const int np = db->n_kf;
    if ((rc = api_ensure_pairs(c, np)) != RGBDFE_OK)
        return rc;
    const uint8_t* dq_desc;
    const rgbdfe_point3f* dq_xyz;
    if ((rc = stage_queries(c, c->verify, q_desc, q_xyz, (size_t)nq, &dq_desc, &dq_xyz)) != RGBDFE_OK)
        return rc;
    rgbdfe_se3_result* d_res = (rgbdfe_se3_result*)results_dev;
    if (!d_res) {
        VerifySlot& sl = c->verify->slot[0];
        if (sl.pending) {
            c->set_error("verify_batch: result slot 0 holds an uncollected rgbdfe_verify_sharded_enqueue");
            return RGBDFE_E_STATE;
        }
        if ((rc = ensure_slot(c, sl, (size_t)np, (size_t)np)) != RGBDFE_OK)
            return rc;
        d_res = sl.d_block;
    }
    const rgbdfe_verify_params vp = { ratio, min_matches, iterations, min_inlier_th, max_mahal, sample_size, seed };
    if ((rc = verify_core(c, db, dq_desc, dq_xyz, nq, 1, np, vp, seed, first_pair_index, 0, d_res)) != RGBDFE_OK)
        return rc;
    if (results) {
        CUDA_TRY(c, cudaMemcpyAsync(results, d_res, (size_t)np * sizeof(rgbdfe_se3_result), cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    }
    return RGBDFE_OK;
}

void rgbdfe_shard_range(int n_kf_total, int rank, int n_ranks, int* lo, int* hi)
{
    if (n_ranks < 1)
        n_ranks = 1;
    const int base = n_kf_total / n_ranks, rem = n_kf_total % n_ranks;
    const int l = rank * base + std::min(rank, rem);
    if (lo)
        *lo = l;
    if (hi)
        *hi = l + base + (rank < rem ? 1 : 0);
}

int rgbdfe_gathered_to_keyframe_order(const rgbdfe_se3_result* gathered, int n_ranks, int n_queries, int n_kf_total, rgbdfe_se3_result* out)
{
    if (!gathered || !out || n_ranks < 1 || n_queries < 1 || n_kf_total < 1)
        return RGBDFE_E_ARG;
    const int block_max = (n_kf_total + n_ranks - 1) / n_ranks;
    for (int r = 0; r < n_ranks; r++) {
        int lo, hi;
        rgbdfe_shard_range(n_kf_total, r, n_ranks, &lo, &hi);
        for (int q = 0; q < n_queries; q++)
            memcpy(out + (size_t)q * n_kf_total + lo, gathered + ((size_t)r * n_queries + q) * block_max, (size_t)(hi - lo) * sizeof(rgbdfe_se3_result));
    }
    return RGBDFE_OK;
}

int rgbdfe_nccl_unique_id(void* id128)
{
    std::string why;
    const NcclApi* n = nccl_api(&why);
    if (!n || !id128)
        return !n ? RGBDFE_E_STATE : RGBDFE_E_ARG;
    NcclId id;
    if (n->GetUniqueId(&id) != 0)
        return RGBDFE_E_CUDA;
    memcpy(id128, &id, sizeof(id));
    return RGBDFE_OK;
}

int rgbdfe_nccl_comm_create(rgbdfe_ctx* c, int n_ranks, int rank, const void* id128, void** comm)
{
    if (!c || !id128 || !comm || n_ranks < 1 || rank < 0 || rank >= n_ranks)
        return RGBDFE_E_ARG;
    std::string why;
    const NcclApi* n = nccl_api(&why);
    if (!n) {
        c->set_error(why);
        return RGBDFE_E_STATE;
    }
    cudaSetDevice(c->device);
    NcclId id;
    memcpy(&id, id128, sizeof(id));
    void* cm = nullptr;
    const int e = n->CommInitRank(&cm, n_ranks, id, rank);
    if (e != 0) {
        c->set_error(std::string("ncclCommInitRank: ") + n->GetErrorString(e));
        return RGBDFE_E_CUDA;
    }
    *comm = cm;
    return RGBDFE_OK;
}

int rgbdfe_nccl_comm_destroy(void* comm)
{
    std::string why;
    const NcclApi* n = nccl_api(&why);
    if (!n || !comm)
        return RGBDFE_E_ARG;
    return n->CommDestroy(comm) == 0 ? RGBDFE_OK : RGBDFE_E_CUDA;
}

int rgbdfe_verify_sharded_enqueue(rgbdfe_ctx* c, rgbdfe_kfdb* db, void* nccl_comm, int n_ranks, int rank, int n_kf_total, const uint8_t* q_desc,
    const rgbdfe_point3f* q_xyz, int nq, int n_queries, const rgbdfe_verify_params* vp, int slot)
{
    int rc = check_query_args(c, db, q_desc, q_xyz, nq);
    if (rc != RGBDFE_OK)
        return rc;
    if (!vp || n_ranks < 1 || rank < 0 || rank >= n_ranks || n_queries < 1 || slot < 0 || slot >= VERIFY_SLOTS || (n_ranks > 1 && !nccl_comm)) {
        c->set_error("verify_sharded: bad argument (ranks, queries, slot in [0, 4), communicator)");
        return RGBDFE_E_ARG;
    }
    int lo, hi;
    rgbdfe_shard_range(n_kf_total, rank, n_ranks, &lo, &hi);
    if (db->n_kf != hi - lo) {
        c->set_error("verify_sharded: this rank's block must hold keyframes [" + std::to_string(lo) + ", " + std::to_string(hi) + ") of "
            + std::to_string(n_kf_total) + " (rgbdfe_shard_range), it holds " + std::to_string(db->n_kf));
        return RGBDFE_E_ARG;
    }
    cudaSetDevice(c->device);
    if ((rc = require_latch(c)) != RGBDFE_OK || (rc = ensure_state(c)) != RGBDFE_OK)
        return rc;
    VerifyState* vs = c->verify;
    VerifySlot& sl = vs->slot[slot];
    if (sl.pending) {
        c->set_error("verify_sharded: slot still holds uncollected results (rgbdfe_verify_sharded_collect first)");
        return RGBDFE_E_STATE;
    }
    const NcclApi* nccl = nullptr;
    if (n_ranks > 1) {
        std::string why;
        if (!(nccl = nccl_api(&why))) {
            c->set_error(why);
            return RGBDFE_E_STATE;
        }
    }
    const int block_max = (n_kf_total + n_ranks - 1) / n_ranks;
    const size_t block_records = (size_t)n_queries * block_max, all_records = block_records * n_ranks;
    const bool fresh = block_records > sl.cap_block || all_records > sl.cap_all;
    const int group = std::max(1, std::min(n_queries, VERIFY_GROUP_PAIRS / block_max));
    if ((rc = api_ensure_pairs(c, group * block_max)) != RGBDFE_OK || (rc = ensure_slot(c, sl, block_records, all_records)) != RGBDFE_OK)
        return rc;
    cudaStream_t s = c->stream, cs = vs->comm_stream;
    if (sl.used)
        CUDA_TRY(c, cudaStreamWaitEvent(s, sl.done, 0)); // the previous gather out of this slot's block has finished
    if (fresh || sl.block_max != block_max || sl.n_queries != n_queries) // padding records of a short block stay zero
        CUDA_TRY(c, cudaMemsetAsync(sl.d_block, 0, block_records * sizeof(rgbdfe_se3_result), s));
    const uint8_t* dq_desc;
    const rgbdfe_point3f* dq_xyz;
    if ((rc = stage_queries(c, vs, q_desc, q_xyz, (size_t)n_queries * nq, &dq_desc, &dq_xyz)) != RGBDFE_OK)
        return rc;
    // pair (q, k): srand(seed + q * n_kf_total + global keyframe index) -- independent of the sharding.
    // Queries go through in groups of about VERIFY_GROUP_PAIRS pairs: one pair table / kNN / prepare / hypotheses pass per group.
    for (int q0 = 0; q0 < n_queries; q0 += group) {
        const int g = std::min(group, n_queries - q0);
        const unsigned seed_q = vp->seed + (unsigned)q0 * (unsigned)n_kf_total;
        if ((rc = verify_core(c, db, dq_desc + (size_t)q0 * nq * 32, dq_xyz + (size_t)q0 * nq, nq, g, block_max, *vp, seed_q, lo, n_kf_total,
                 sl.d_block + (size_t)q0 * block_max))
            != RGBDFE_OK)
            return rc;
    }
    CUDA_TRY(c, cudaEventRecord(sl.computed, s));
    CUDA_TRY(c, cudaStreamWaitEvent(cs, sl.computed, 0));
    const rgbdfe_se3_result* src = sl.d_block;
    if (n_ranks > 1) {
        const int e = nccl->AllGather(sl.d_block, sl.d_all, block_records * sizeof(rgbdfe_se3_result), 0 /* ncclInt8 */, nccl_comm, cs);
        if (e != 0) {
            c->set_error(std::string("ncclAllGather: ") + nccl->GetErrorString(e));
            return RGBDFE_E_CUDA;
        }
        src = sl.d_all;
    }
    CUDA_TRY(c, cudaMemcpyAsync(sl.h_all, src, all_records * sizeof(rgbdfe_se3_result), cudaMemcpyDeviceToHost, cs));
    CUDA_TRY(c, cudaEventRecord(sl.done, cs));
    sl.pending = sl.used = true;
    sl.n_ranks = n_ranks, sl.n_queries = n_queries, sl.block_max = block_max, sl.n_kf_total = n_kf_total;
    return RGBDFE_OK;
}

int rgbdfe_verify_sharded_collect(rgbdfe_ctx* c, int slot, rgbdfe_se3_result* results)
{
    if (!c || slot < 0 || slot >= VERIFY_SLOTS)
        return RGBDFE_E_ARG;
    if (!c->verify || !c->verify->slot[slot].pending) {
        c->set_error("verify_sharded_collect: nothing enqueued on this slot");
        return RGBDFE_E_STATE;
    }
    VerifySlot& sl = c->verify->slot[slot];
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaEventSynchronize(sl.done));
    sl.pending = false;
    if (results)
        return rgbdfe_gathered_to_keyframe_order(sl.h_all, sl.n_ranks, sl.n_queries, sl.n_kf_total, results);
    return RGBDFE_OK;
}

int rgbdfe_verify_sharded(rgbdfe_ctx* c, rgbdfe_kfdb* db, void* nccl_comm, int n_ranks, int rank, int n_kf_total, const uint8_t* q_desc,
    const rgbdfe_point3f* q_xyz, int nq, int n_queries, const rgbdfe_verify_params* vp, rgbdfe_se3_result* results)
{
    const int rc = rgbdfe_verify_sharded_enqueue(c, db, nccl_comm, n_ranks, rank, n_kf_total, q_desc, q_xyz, nq, n_queries, vp, 0);
    if (rc != RGBDFE_OK)
        return rc;
    return rgbdfe_verify_sharded_collect(c, 0, results);
}

} // extern "C"
