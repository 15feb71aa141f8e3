// Row N4 -- LoopDetector::add / LoopDetector::obtainCandidates (PlaceRecognition/LoopDetector.cpp:22-84), the
// candidate generation in front of the loop-pair verifier (PoseGraph::detectLoop, Solver/PoseGraph.cpp:254).
//
// The bag-of-words vectors themselves come from DBoW3 (third party, not under /root/reference, version unpinned in
// CMakeLists.txt): Frame::computeBoW -> Vocabulary::transform (Core/Frame.cpp:243-249).  They are INPUTS here: a
// DBoW3::BowVector is a std::map<WordId, WordValue>, i.e. (word id ascending, double weight) pairs.  What this file
// computes is everything obtainCandidates does with them:
//   * Vocabulary::score for the ORB vocabulary's L1_NORM scoring (DBoW3 ScoringObject.cpp, L1Scoring::score): over the
//     words both vectors hold, in ascending word order, s += |v - w| - |v| - |w| in double; score = -s / 2;
//   * the minimum score over the connected keyframes (:38-47; argument order score(KFi, KF) kept: the two
//     subtractions are not symmetric in floating point);
//   * the inverted-file walk (:50-64): keyframes sharing at least one word, in order of first encounter (query words
//     ascending, each word's list in insertion order), skipping the query, its connected keyframes and keyframes whose
//     mnSharingWord mark already equals the query id (the mark persists between calls, as the member does);
//   * score > minScore && |id - id_q| > interval (:69-76), and when more than five remain std::sort by
//     Frame::CompareBoWscore and the first five (:78-81) -- libstdc++'s introsort permutation (introsort.cuh), so that
//     ties resolve as they do on the host.
// One thread per keyframe merges its word list with the query's (k_bow_scores); one CTA then replays the list logic
// (k_loop_select): encounter order = ascending (first shared word, insertion index) by a bitonic sort of 64-bit keys.
#include <algorithm>
#include <climits>
#include <new>

#include "common.cuh"
#include "introsort.cuh"

using namespace rgbdfe;

struct rgbdfe_bowdb {
    int device = 0;
    int max_kf = 0, n_kf = 0;
    size_t max_nnz = 0, nnz = 0;
    int* off = nullptr;      // [max_kf + 1] CSR offsets
    int* word = nullptr;     // [max_nnz] ascending per keyframe
    double* weight = nullptr;
    int* kf_id = nullptr;    // Frame::id()
    int* mark = nullptr;     // Frame::mnSharingWord (uninitialised in the reference; -1 here)
    int* first_word = nullptr; // per-call scratch: smallest shared word or INT_MAX
    double* score_qk = nullptr; // score(query, KF_k)
    double* score_kq = nullptr; // score(KF_k, query)
    int* d_conn = nullptr;   // connected indices of a call
    int* d_out = nullptr;    // [0] = n, then indices; scores behind
    double* d_out_score = nullptr;
    int* h_out = nullptr;    // pinned: n + indices (64 ints), then 64 doubles
    std::vector<int> h_off;  // host copy of the offsets (append bookkeeping)
};

namespace {

template <class T>
cudaError_t dmalloc(T** p, size_t n)
{
    return cudaMalloc((void**)p, n * sizeof(T) > 0 ? n * sizeof(T) : 1);
}

// DBoW3 L1Scoring::score(v1, v2) accumulated over the common words in ascending order; returns the raw sum
// (the caller applies -s / 2) and the first common word.
__device__ __forceinline__ void merge_scores(const int* __restrict__ w1, const double* __restrict__ x1, int n1, const int* __restrict__ w2,
    const double* __restrict__ x2, int n2, double& s12, double& s21, int& first)
{
    int i = 0, j = 0;
    double a = 0.0, b = 0.0;
    first = INT_MAX;
    while (i < n1 && j < n2) {
        const int p = w1[i], q = w2[j];
        if (p == q) {
            const double vi = x1[i], wi = x2[j];
            a += fabs(vi - wi) - fabs(vi) - fabs(wi); // score(v1, v2)
            b += fabs(wi - vi) - fabs(wi) - fabs(vi); // score(v2, v1)
            if (first == INT_MAX)
                first = p;
            i++, j++;
        } else if (p < q)
            i++;
        else
            j++;
    }
    s12 = a, s21 = b;
}

__global__ void k_bow_scores(const int* __restrict__ off, const int* __restrict__ word, const double* __restrict__ weight, int n_kf, int q,
    int* __restrict__ first_word, double* __restrict__ score_qk, double* __restrict__ score_kq)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_kf)
        return;
    const int qo = off[q], qn = off[q + 1] - qo, ko = off[k], kn = off[k + 1] - ko;
    double s_qk, s_kq;
    int first;
    merge_scores(word + qo, weight + qo, qn, word + ko, weight + ko, kn, s_qk, s_kq, first);
    first_word[k] = first;
    score_qk[k] = -s_qk / 2.0;
    score_kq[k] = -s_kq / 2.0;
}

struct ScoreIdx {
    double s;
    int k;
};
struct ScoreGreater { // Frame::CompareBoWscore (Core/Frame.h:103)
    __host__ __device__ bool operator()(const ScoreIdx& a, const ScoreIdx& b) const { return a.s > b.s; }
};

// dynamic smem: unsigned long long keys[npow2]; unsigned char conn[n_kf]
__global__ void __launch_bounds__(1024) k_loop_select(int n_kf, int npow2, int q, const int* __restrict__ conn_list, int n_conn, int interval,
    const int* __restrict__ kf_id, int* __restrict__ mark, const int* __restrict__ first_word, const double* __restrict__ score_qk,
    const double* __restrict__ score_kq, ScoreIdx* __restrict__ list /* [n_kf] global scratch */, int* __restrict__ out, double* __restrict__ out_score,
    int cap)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem);
    unsigned char* conn = smem + (size_t)npow2 * 8;
    __shared__ double s_min[32];
    __shared__ int s_scan[32];
    __shared__ int s_total;
    const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5;
    const int qid = kf_id[q];
    for (int i = tid; i < n_kf; i += nthr)
        conn[i] = 0;
    __syncthreads();
    for (int i = tid; i < n_conn; i += nthr)
        conn[conn_list[i]] = 1;
    __syncthreads();
    // minimum score with the connected keyframes (:38-47)
    double mn = 1.7976931348623157e308; // numeric_limits<double>::max()
    for (int i = tid; i < n_kf; i += nthr)
        if (conn[i] && i != q)
            mn = fmin(mn, score_kq[i]);
    for (int o = 16; o > 0; o >>= 1)
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    if (lane == 0)
        s_min[warp] = mn;
    __syncthreads();
    mn = s_min[0];
    for (int w = 1; w < (nthr >> 5); w++)
        mn = fmin(mn, s_min[w]);
    const double minScore = mn;
    // keyframes sharing a word (:50-64); key = (first shared word, insertion index) = order of first encounter
    for (int i = tid; i < npow2; i += nthr) {
        unsigned long long key = ~0ull;
        if (i < n_kf && i != q && first_word[i] != INT_MAX && mark[i] != qid && !conn[i]) {
            mark[i] = qid;
            key = ((unsigned long long)(unsigned)first_word[i] << 32) | (unsigned)i;
        }
        keys[i] = key;
    }
    __syncthreads();
    for (int k = 2; k <= npow2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < npow2; i += nthr) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long a = keys[i], b = keys[ixj];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[ixj] = a;
                    }
                }
            }
            __syncthreads();
        }
    // possibleLoops in encounter order (:69-76): ordered compaction
    if (tid == 0)
        s_total = 0;
    __syncthreads();
    for (int i0 = 0; i0 < npow2; i0 += nthr) {
        const int i = i0 + tid;
        bool keep = false;
        int k = 0;
        if (i < npow2 && keys[i] != ~0ull) {
            k = (int)(keys[i] & 0xffffffffu);
            const int d = kf_id[k] - qid;
            keep = score_qk[k] > minScore && (d < 0 ? -d : d) > interval;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0)
            s_scan[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
        for (int w = 0; w < (nthr >> 5); w++) {
            before += w < warp ? s_scan[w] : 0;
            total += s_scan[w];
        }
        const int base = s_total;
        if (keep) {
            ScoreIdx e;
            e.s = score_qk[k];
            e.k = k;
            list[base + before + __popc(bal & ((1u << lane) - 1))] = e;
        }
        __syncthreads();
        if (tid == 0)
            s_total = base + total;
        __syncthreads();
    }
    if (tid == 0) {
        int n = s_total;
        if (n > 5) { // :78-81
            std_sort(list, list + n, ScoreGreater());
            n = 5;
        }
        n = min(n, cap);
        out[0] = n;
        for (int i = 0; i < n; i++) {
            out[1 + i] = list[i].k;
            out_score[i] = list[i].s;
        }
    }
}

} // namespace

extern "C" {

int rgbdfe_bowdb_create(rgbdfe_ctx* c, rgbdfe_bowdb** out, int max_kf, int max_words_per_kf)
{
    if (!c || !out || max_kf < 1 || max_words_per_kf < 1)
        return RGBDFE_E_ARG;
    *out = nullptr;
    if (max_kf > 16384) {
        c->set_error("bowdb: at most 16384 keyframes (the candidate ordering runs in one CTA's shared memory)");
        return RGBDFE_E_CAPACITY;
    }
    cudaSetDevice(c->device);
    rgbdfe_bowdb* db = new (std::nothrow) rgbdfe_bowdb();
    if (!db)
        return RGBDFE_E_ARG;
    db->device = c->device, db->max_kf = max_kf, db->max_nnz = (size_t)max_kf * max_words_per_kf;
    cudaError_t e = dmalloc(&db->off, (size_t)max_kf + 1);
    auto chain = [&](cudaError_t r) {
        if (e == cudaSuccess)
            e = r;
    };
    chain(dmalloc(&db->word, db->max_nnz));
    chain(dmalloc(&db->weight, db->max_nnz));
    chain(dmalloc(&db->kf_id, (size_t)max_kf));
    chain(dmalloc(&db->mark, (size_t)max_kf));
    chain(dmalloc(&db->first_word, (size_t)max_kf));
    chain(dmalloc(&db->score_qk, (size_t)max_kf));
    chain(dmalloc(&db->score_kq, (size_t)max_kf));
    chain(dmalloc(&db->d_conn, (size_t)max_kf));
    chain(dmalloc(&db->d_out, 64));
    chain(dmalloc(&db->d_out_score, (size_t)max_kf * 2 + 64)); // [0, 64) scores out; the rest is the ScoreIdx list (16 B per keyframe)
    chain(cudaMallocHost((void**)&db->h_out, 64 * sizeof(int) + 64 * sizeof(double)));
    if (e == cudaSuccess)
        e = cudaMemset(db->mark, 0xff, (size_t)max_kf * sizeof(int));
    if (e == cudaSuccess)
        e = cudaMemset(db->off, 0, sizeof(int));
    if (e != cudaSuccess) {
        c->set_error(std::string("bowdb_create: ") + cudaGetErrorString(e));
        rgbdfe_bowdb_destroy(c, db);
        return RGBDFE_E_CUDA;
    }
    db->h_off.assign(1, 0);
    *out = db;
    return RGBDFE_OK;
}

void rgbdfe_bowdb_destroy(rgbdfe_ctx* c, rgbdfe_bowdb* db)
{
    if (!db)
        return;
    if (c) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
    }
    cudaFree(db->off), cudaFree(db->word), cudaFree(db->weight), cudaFree(db->kf_id), cudaFree(db->mark), cudaFree(db->first_word);
    cudaFree(db->score_qk), cudaFree(db->score_kq), cudaFree(db->d_conn), cudaFree(db->d_out), cudaFree(db->d_out_score);
    cudaFreeHost(db->h_out);
    delete db;
}

int rgbdfe_bowdb_add(rgbdfe_ctx* c, rgbdfe_bowdb* db, int kf_id, const int* words, const double* weights, int n, int* index)
{
    if (!c || !db || n < 0 || (n > 0 && (!words || !weights)))
        return RGBDFE_E_ARG;
    if (db->n_kf >= db->max_kf || db->nnz + (size_t)n > db->max_nnz) {
        c->set_error("bowdb_add: database full");
        return RGBDFE_E_CAPACITY;
    }
    for (int i = 1; i < n; i++)
        if (words[i] <= words[i - 1]) {
            c->set_error("bowdb_add: word ids must be strictly ascending (a DBoW3::BowVector is a std::map)");
            return RGBDFE_E_ARG;
        }
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    const int k = db->n_kf;
    const int end = (int)(db->nnz + (size_t)n);
    if (n > 0) {
        CUDA_TRY(c, cudaMemcpyAsync(db->word + db->nnz, words, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
        CUDA_TRY(c, cudaMemcpyAsync(db->weight + db->nnz, weights, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    db->h_off.push_back(end);
    CUDA_TRY(c, cudaMemcpyAsync(db->off + k + 1, &db->h_off[k + 1], sizeof(int), cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(db->kf_id + k, &kf_id, sizeof(int), cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaStreamSynchronize(s)); // the sources are caller / stack memory
    db->nnz = (size_t)end;
    db->n_kf = k + 1;
    if (index)
        *index = k;
    return RGBDFE_OK;
}

int rgbdfe_loop_candidates(rgbdfe_ctx* c, rgbdfe_bowdb* db, int query_index, const int* connected, int n_connected, int interval, int* candidates,
    double* scores, int cap, int* n)
{
    if (!c || !db || !n || query_index < 0 || query_index >= db->n_kf || n_connected < 0 || (n_connected > 0 && !connected) || cap < 0
        || (cap > 0 && !candidates))
        return RGBDFE_E_ARG;
    *n = 0;
    if (db->device != c->device) {
        c->set_error("loop_candidates: the database lives on another device than the context");
        return RGBDFE_E_ARG;
    }
    if (n_connected == 0) // :33-35
        return RGBDFE_OK;
    for (int i = 0; i < n_connected; i++)
        if (connected[i] < 0 || connected[i] >= db->n_kf) {
            c->set_error("loop_candidates: connected keyframe index out of range");
            return RGBDFE_E_ARG;
        }
    if (n_connected > db->max_kf)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    CUDA_TRY(c, cudaMemcpyAsync(db->d_conn, connected, (size_t)n_connected * sizeof(int), cudaMemcpyHostToDevice, s));
    const int n_kf = db->n_kf;
    k_bow_scores<<<(n_kf + 127) / 128, 128, 0, s>>>(db->off, db->word, db->weight, n_kf, query_index, db->first_word, db->score_qk, db->score_kq);
    int npow2 = 32;
    while (npow2 < n_kf)
        npow2 <<= 1;
    const size_t smem = (size_t)npow2 * 8 + (size_t)n_kf;
    CUDA_TRY(c, ensure_dyn_smem(c, k_loop_select, smem));
    k_loop_select<<<1, 1024, smem, s>>>(n_kf, npow2, query_index, db->d_conn, n_connected, interval, db->kf_id, db->mark, db->first_word, db->score_qk,
        db->score_kq, reinterpret_cast<ScoreIdx*>(db->d_out_score + 64), db->d_out, db->d_out_score, std::min(cap, 5));
    c->launches += 2;
    CUDA_TRY(c, cudaGetLastError());
    double* h_sc = reinterpret_cast<double*>(db->h_out + 64);
    CUDA_TRY(c, cudaMemcpyAsync(db->h_out, db->d_out, 8 * sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaMemcpyAsync(h_sc, db->d_out_score, 8 * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    const int cnt = db->h_out[0];
    *n = cnt;
    for (int i = 0; i < cnt; i++) {
        candidates[i] = db->h_out[1 + i];
        if (scores)
            scores[i] = h_sc[i];
    }
    return RGBDFE_OK;
}

} // extern "C"
