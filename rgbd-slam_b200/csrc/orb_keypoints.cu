// Keypoint-side ORB kernels:
//   K3 k_quadtree — ORBextractor::DistributeOctTree + ExtractorNode::divideNode
//        (Features/ORBextractor.cpp:414-611, Features/ExtractorNode.cpp:5-57), one CTA per
//        (frame, level).  Label-based, pass-synchronous formulation: points never move,
//        each carries the list position of its node; a pass = quadrant histogram (smem
//        atomics) -> choose the nodes to split (all in the main phase; the largest-first
//        prefix that reaches N in the final phase: bitonic sort + scan) -> rebuild the
//        ordered node list with exclusive scans -> relabel.  "first key with the maximum
//        response" == (max response, min candidate index), so no per-node order is kept.
//        Prologue: ordered compaction of the per-cell FAST lists (cell-major).
//   K4+K6 k_orient_describe — IC_Angle (:16-41) + computeOrbDescriptor (:45-87) + keypoint
//        rescale / concatenation (:755-764) + Frame::uprojectCamera (Core/Frame.cpp:91-117),
//        one warp per keypoint.  FP32 without FMA contraction (-fmad=false); sin/cos restate
//        glibc's sinf/cosf (double arithmetic, rounded once) so descriptors are bit-exact.
#include "brief_pattern.h"
#include "common.cuh"
#include "frame_tail.cuh"
#include "orb_math.cuh"

namespace rgbdfe {

// bit_pattern_31_ as int8 (x0, y0, x1, y1) x 256.  Lane L of a warp reads bytes [32 L, 32 L + 32): a per-lane index,
// which constant memory would serialise (one address per cycle through the ADU), so the table lives in global
// memory and is staged into shared memory by every CTA.
__device__ __align__(16) signed char g_pattern[1024];

struct QtNode {
    short ulx, uly, brx, bry;
    int count;
    int seq;
};

__device__ int block_exscan(int* a, int len, int* partial)
{
    const int tid = threadIdx.x;
    const int chunk = (len + QT_THREADS - 1) / QT_THREADS;
    const int b = min(tid * chunk, len), e = min(b + chunk, len);
    int s = 0;
    for (int i = b; i < e; i++)
        s += a[i];
    partial[tid] = s;
    __syncthreads();
    for (int off = 1; off < QT_THREADS; off <<= 1) {
        int v = tid >= off ? partial[tid - off] : 0;
        __syncthreads();
        partial[tid] += v;
        __syncthreads();
    }
    int base = tid ? partial[tid - 1] : 0;
    const int total = partial[QT_THREADS - 1];
    __syncthreads();
    for (int i = b; i < e; i++) {
        int v = a[i];
        a[i] = base;
        base += v;
    }
    __syncthreads();
    return total;
}

__device__ __forceinline__ int quadrant_of(const QtNode& q, unsigned c)
{
    const int halfX = (int)ceilf(static_cast<float>(q.brx - q.ulx) / 2);
    const int halfY = (int)ceilf(static_cast<float>(q.bry - q.uly) / 2);
    const int right = !(cand_x(c) < q.ulx + halfX);
    const int down = !(cand_y(c) < q.uly + halfY);
    return right + 2 * down; // n1=0 n2=1 n3=2 n4=3
}

// dynamic smem layout (NC = node capacity, SC = pow2 >= NC):
//   QtNode A[NC], B[NC]; int child[NC*4]; int newpos[NC]; int isproc[NC]; int scan[NC];
//   int order[NC]; unsigned best[NC]; unsigned long long keys[SC]; int partial[QT_THREADS]
__global__ void __launch_bounds__(QT_THREADS) k_quadtree(const OrbConfig* __restrict__ cfg, const int* __restrict__ cell_count,
    const unsigned* __restrict__ cell_list, int cell_slot, unsigned* __restrict__ cand_all, int* __restrict__ label_all, int cand_slot,
    int* __restrict__ cand_count, SelKp* __restrict__ sel_all, int sel_slot, int* __restrict__ sel_count, int NC, int SC)
{
    const int level = blockIdx.x, slot = blockIdx.y, tid = threadIdx.x;
    const LevelDesc L = cfg->lv[level];
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem);
    QtNode* nodeA = reinterpret_cast<QtNode*>(keys + SC);
    QtNode* nodeB = nodeA + NC;
    int* child = reinterpret_cast<int*>(nodeB + NC);
    int* newpos = child + NC * 4;
    int* isproc = newpos + NC;
    int* scan = isproc + NC;
    int* order = scan + NC;
    unsigned* best = reinterpret_cast<unsigned*>(order + NC);
    int* partial = reinterpret_cast<int*>(best + NC);
    __shared__ int s_L, s_finish, s_phase, s_K, s_nToExpand, s_seq, s_overflow;

    unsigned* cand = cand_all + (size_t)slot * cand_slot + L.cand_off;
    int* label = label_all + (size_t)slot * cand_slot + L.cand_off;
    SelKp* sel = sel_all + (size_t)slot * sel_slot + L.sel_off;

    // ---- prologue: cell-major ordered compaction of the FAST lists ----
    const int ncells = L.nCols * L.nRows;
    const int* ccount = cell_count + (size_t)slot * cell_slot + L.cell_off;
    const unsigned* clist = cell_list + ((size_t)slot * cell_slot + L.cell_off) * CELL_CAP;
    int carry = 0;
    for (int base = 0; base < ncells; base += QT_THREADS) {
        int cidx = base + tid;
        int v = cidx < ncells ? ccount[cidx] : 0;
        partial[tid] = v;
        __syncthreads();
        for (int off = 1; off < QT_THREADS; off <<= 1) {
            int t = tid >= off ? partial[tid - off] : 0;
            __syncthreads();
            partial[tid] += t;
            __syncthreads();
        }
        if (cidx < ncells)
            label[cidx] = carry + partial[tid] - v; // temp: exclusive offset of the cell
        carry += partial[QT_THREADS - 1];
        __syncthreads();
    }
    const int n = min(carry, L.cand_cap);
    __syncthreads();
    for (int cidx = tid >> 5; cidx < ncells; cidx += QT_THREADS / 32) {
        const int off = label[cidx], cnt = ccount[cidx];
        for (int k = tid & 31; k < cnt; k += 32)
            if (off + k < L.cand_cap)
                cand[off + k] = clist[(size_t)cidx * CELL_CAP + k];
    }
    __syncthreads();
    if (tid == 0)
        cand_count[slot * RGBDFE_MAX_LEVELS + level] = n;
    const int N = L.quota;
    if (n == 0) {
        if (tid == 0)
            sel_count[slot * RGBDFE_MAX_LEVELS + level] = 0;
        return;
    }

    // ---- initial nodes (ORBextractor.cpp:418-456) ----
    const int minX = 16, maxX = L.maxBorderX, minY = 16, maxY = L.maxBorderY;
    int nIni = (int)roundf(static_cast<float>(maxX - minX) / (maxY - minY));
    nIni = max(nIni, 1); // the reference divides by zero here for very tall images
    const float hX = static_cast<float>(maxX - minX) / nIni;
    QtNode* cur = nodeA;
    QtNode* nxt = nodeB;
    for (int i = tid; i < nIni && i < NC; i += QT_THREADS) {
        QtNode q;
        q.ulx = (short)(int)(hX * static_cast<float>(i));
        q.brx = (short)(int)(hX * static_cast<float>(i + 1));
        q.uly = 0;
        q.bry = (short)(maxY - minY);
        q.count = 0;
        q.seq = i;
        nxt[i] = q;
    }
    __syncthreads();
    for (int p = tid; p < n; p += QT_THREADS) {
        int k = (int)((float)cand_x(cand[p]) / hX);
        k = min(k, nIni - 1);
        label[p] = k;
        atomicAdd(&nxt[k].count, 1);
    }
    __syncthreads();
    if (tid == 0) {
        int Lc = 0;
        for (int i = 0; i < nIni; i++) {
            newpos[i] = Lc;
            if (nxt[i].count > 0)
                cur[Lc++] = nxt[i];
        }
        s_L = Lc;
        s_finish = 0;
        s_phase = 0;
        s_seq = nIni;
        s_overflow = 0;
    }
    __syncthreads();
    for (int p = tid; p < n; p += QT_THREADS)
        label[p] = newpos[label[p]];
    __syncthreads();

    // ---- passes ----
    while (!s_finish) {
        const int Lc = s_L;
        const int phase = s_phase;
        for (int i = tid; i < Lc * 4; i += QT_THREADS)
            child[i] = 0;
        if (tid == 0)
            s_nToExpand = 0;
        __syncthreads();
        for (int p = tid; p < n; p += QT_THREADS) {
            const int id = label[p];
            const QtNode q = cur[id];
            if (q.count > 1)
                atomicAdd(&child[id * 4 + quadrant_of(q, cand[p])], 1);
        }
        __syncthreads();
        int K;
        if (phase == 0) {
            // main phase: every expandable node, in list order
            for (int i = tid; i < Lc; i += QT_THREADS)
                isproc[i] = scan[i] = cur[i].count > 1;
            __syncthreads();
            K = block_exscan(scan, Lc, partial);
            for (int i = tid; i < Lc; i += QT_THREADS)
                if (isproc[i])
                    order[scan[i]] = i;
            __syncthreads();
        } else {
            // final phase: largest (size, seq) first until the list reaches N nodes
            for (int i = tid; i < SC; i += QT_THREADS) {
                unsigned long long k = 0;
                if (i < Lc && cur[i].count > 1)
                    k = ((unsigned long long)cur[i].count << 40) | ((unsigned long long)cur[i].seq << 16) | (unsigned long long)i;
                keys[i] = k;
            }
            __syncthreads();
            for (int size = 2; size <= SC; size <<= 1) {
                for (int stride = size >> 1; stride > 0; stride >>= 1) {
                    for (int t = tid; t < SC / 2; t += QT_THREADS) {
                        int lo = 2 * t - (t & (stride - 1));
                        int hi = lo + stride;
                        bool desc = (lo & size) == 0;
                        unsigned long long a = keys[lo], b = keys[hi];
                        if ((a < b) == desc) {
                            keys[lo] = b;
                            keys[hi] = a;
                        }
                    }
                    __syncthreads();
                }
            }
            // m = number of expandable nodes; scan (nonempty children - 1) in sorted order
            for (int j = tid; j < Lc; j += QT_THREADS) {
                int v = 0;
                if (keys[j] != 0) {
                    const int i = (int)(keys[j] & 0xffff);
                    v = (child[i * 4] > 0) + (child[i * 4 + 1] > 0) + (child[i * 4 + 2] > 0) + (child[i * 4 + 3] > 0) - 1;
                }
                scan[j] = v;
                isproc[j] = 0;
            }
            if (tid == 0)
                s_K = -1;
            __syncthreads();
            // keep the per-element values: inclusive = exclusive + v
            for (int j = tid; j < Lc; j += QT_THREADS)
                newpos[j] = scan[j];
            __syncthreads();
            block_exscan(scan, Lc, partial);
            for (int j = tid; j < Lc; j += QT_THREADS) {
                if (keys[j] == 0)
                    continue;
                const int incl = Lc + scan[j] + newpos[j];
                const int before = Lc + scan[j];
                if (incl >= N && (j == 0 || before < N))
                    atomicMax(&s_K, j + 1);
            }
            __syncthreads();
            if (tid == 0 && s_K < 0) {
                int m = 0; // all expandable nodes are processed
                for (int j = 0; j < Lc; j++)
                    m += keys[j] != 0;
                s_K = m;
            }
            __syncthreads();
            K = s_K;
            for (int t = tid; t < K; t += QT_THREADS) {
                const int i = (int)(keys[t] & 0xffff);
                order[t] = i;
                isproc[i] = 1;
            }
            __syncthreads();
        }
        // children per processed node -> positions and creation sequence numbers
        for (int t = tid; t < K; t += QT_THREADS) {
            const int i = order[t];
            scan[t] = (child[i * 4] > 0) + (child[i * 4 + 1] > 0) + (child[i * 4 + 2] > 0) + (child[i * 4 + 3] > 0);
        }
        __syncthreads();
        // cnt copy (exscan overwrites): stash in best[] temporarily
        for (int t = tid; t < K; t += QT_THREADS)
            best[t] = (unsigned)scan[t];
        __syncthreads();
        const int C = block_exscan(scan, K, partial); // scan[t] = pre[t]
        // rest positions
        for (int i = tid; i < Lc; i += QT_THREADS)
            newpos[i] = !isproc[i];
        __syncthreads();
        const int R = block_exscan(newpos, Lc, partial);
        const int newL = C + R;
        if (newL > NC) { // cannot happen for sane quotas; refuse to corrupt memory
            if (tid == 0) {
                s_overflow = 1;
                s_finish = 1;
            }
            __syncthreads();
            break;
        }
        const int seqBase = s_seq;
        for (int t = tid; t < K; t += QT_THREADS) {
            const int i = order[t];
            const QtNode q = cur[i];
            const int halfX = (int)ceilf(static_cast<float>(q.brx - q.ulx) / 2);
            const int halfY = (int)ceilf(static_cast<float>(q.bry - q.uly) / 2);
            int cc[4] = { child[i * 4], child[i * 4 + 1], child[i * 4 + 2], child[i * 4 + 3] };
            const int cnt = (int)best[t];
            const int start = C - scan[t] - cnt; // groups appear in reverse processing order
            int created = 0, expandable = 0;
            for (int k = 0; k < 4; k++) {
                if (cc[k] == 0) {
                    child[i * 4 + k] = -1;
                    continue;
                }
                QtNode c;
                c.ulx = (k & 1) ? (short)(q.ulx + halfX) : q.ulx;
                c.brx = (k & 1) ? q.brx : (short)(q.ulx + halfX);
                c.uly = (k & 2) ? (short)(q.uly + halfY) : q.uly;
                c.bry = (k & 2) ? q.bry : (short)(q.uly + halfY);
                c.count = cc[k];
                c.seq = seqBase + scan[t] + created;
                const int pos = start + (cnt - 1 - created); // n4 frontmost, n1 last in the group
                nxt[pos] = c;
                child[i * 4 + k] = pos;
                created++;
                expandable += cc[k] > 1;
            }
            if (expandable)
                atomicAdd(&s_nToExpand, expandable);
        }
        for (int i = tid; i < Lc; i += QT_THREADS)
            if (!isproc[i]) {
                newpos[i] += C;
                nxt[newpos[i]] = cur[i];
            }
        __syncthreads();
        for (int p = tid; p < n; p += QT_THREADS) {
            const int id = label[p];
            if (isproc[id])
                label[p] = child[id * 4 + quadrant_of(cur[id], cand[p])];
            else
                label[p] = newpos[id];
        }
        __syncthreads();
        if (tid == 0) {
            s_seq = seqBase + C;
            s_L = newL;
            if (newL >= N || newL == Lc)
                s_finish = 1;
            else if (phase == 0 && newL + s_nToExpand * 3 > N)
                s_phase = 1;
        }
        QtNode* tmp = cur;
        cur = nxt;
        nxt = tmp;
        __syncthreads();
    }

    // ---- best point per node (ORBextractor.cpp:592-608), list order ----
    const int Lf = s_L;
    for (int i = tid; i < Lf; i += QT_THREADS)
        best[i] = 0;
    __syncthreads();
    for (int p = tid; p < n; p += QT_THREADS)
        atomicMax(&best[label[p]], ((unsigned)cand_s(cand[p]) << 24) | (0xffffffu - (unsigned)p));
    __syncthreads();
    const int nout = min(Lf, L.sel_cap);
    for (int i = tid; i < nout; i += QT_THREADS) {
        const unsigned c = cand[0xffffffu - (best[i] & 0xffffffu)];
        SelKp s;
        s.x = (short)(cand_x(c) + 16);
        s.y = (short)(cand_y(c) + 16);
        s.score = cand_s(c);
        s.angle = 0.f;
        sel[i] = s;
    }
    if (tid == 0)
        sel_count[slot * RGBDFE_MAX_LEVELS + level] = s_overflow ? -1 : nout;
}

// ------------------------------------------------------------- K4 + K6 ----
__global__ void __launch_bounds__(256) k_orient_describe(const OrbConfig* __restrict__ cfg, const uint8_t* __restrict__ pyr, size_t pyr_slot,
    const uint8_t* __restrict__ blur, size_t blur_slot, const SelKp* __restrict__ sel_all, int sel_slot, const int* __restrict__ sel_count,
    const FrameBatch frames /* by value: one frame struct per slot */, const __grid_constant__ ImgPtrs depth, int depth_stride, int unproject,
    const __grid_constant__ FramePrep prep)
{
    __shared__ __align__(16) signed char s_pattern[1024];
    reinterpret_cast<int*>(s_pattern)[threadIdx.x] = reinterpret_cast<const int*>(g_pattern)[threadIdx.x]; // 256 threads x 4 B
    __syncthreads();
    const int slot = blockIdx.y;
    const int lane = threadIdx.x & 31;
    int gi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); // output keypoint index
    const int* cnt = sel_count + slot * RGBDFE_MAX_LEVELS;
    int level = 0, base = 0;
    const int nl = cfg->nlevels;
    while (level < nl && gi >= base + max(cnt[level], 0)) {
        base += max(cnt[level], 0);
        level++;
    }
    const rgbdfe_frame F = frames.f[slot];
    if (level >= nl) {
        if (gi == base && lane == 0)
            *F.n = min(base, F.cap); // first warp past the end publishes the total
        return;
    }
    if (gi >= F.cap)
        return;
    const LevelDesc L = cfg->lv[level];
    const SelKp s = sel_all[(size_t)slot * sel_slot + L.sel_off + (gi - base)];

    // ---- IC_Angle: integer moments over the 749-pixel disc ----
    const uint8_t* center = pyr + slot * pyr_slot + L.pyr_off + (size_t)(EDGE_THRESHOLD + s.y) * L.pstride + EDGE_THRESHOLD + s.x;
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int u = lane - 15;
        const int au = abs(u);
        m10 = u * center[u];
        for (int v = 1; v <= 15; v++) {
            if (au <= cfg->umax[v]) {
                int vp = center[u + v * L.pstride], vm = center[u - v * L.pstride];
                m10 += u * (vp + vm);
                m01 += v * (vp - vm);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- steered BRIEF: lane = descriptor byte ----
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    sincosf_glibc(__fmul_rn(angle, factorPI), &b, &a);
    const uint8_t* bc = blur + slot * blur_slot + L.blur_off + (size_t)s.y * L.bstride + s.x;
    signed char pat[32];
    *reinterpret_cast<uint4*>(pat) = *reinterpret_cast<const uint4*>(s_pattern + lane * 32);
    *reinterpret_cast<uint4*>(pat + 16) = *reinterpret_cast<const uint4*>(s_pattern + lane * 32 + 16);
    unsigned val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float x0 = (float)pat[k * 4], y0 = (float)pat[k * 4 + 1], x1 = (float)pat[k * 4 + 2], y1 = (float)pat[k * 4 + 3];
        int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        int t0 = bc[r0 * L.bstride + c0], t1 = bc[r1 * L.bstride + c1];
        val |= (unsigned)(t0 < t1) << k;
    }
    F.desc[(size_t)gi * 32 + lane] = (uint8_t)val;

    if (lane == 0) {
        rgbdfe_keypoint kp;
        float px = (float)s.x, py = (float)s.y;
        if (level != 0) {
            px = __fmul_rn(px, L.scale);
            py = __fmul_rn(py, L.scale);
        }
        kp.x = px;
        kp.y = py;
        kp.size = (float)L.patch_size;
        kp.angle = angle;
        kp.response = (float)s.score;
        kp.octave = level;
        kp.class_id = -1;
        store_frame_keypoint(F, gi, kp, prep, depth.p[slot], depth_stride, unproject);
    }
}

// Frame::uprojectCamera for a depth map that stays in pinned HOST memory (option "host_zero_copy"): one CTA per frame,
// every thread issues its (up to) 8 PCIe sector reads before it uses the first -- ~2000 reads per frame in flight on a
// handful of warps, instead of one read at the end of each K6 warp, which holds that warp's slot for the PCIe round trip.
#define UNPROJ_THREADS 256
#define UNPROJ_ILP 8
__global__ void __launch_bounds__(UNPROJ_THREADS) k_unproject_host(const FrameBatch frames, const __grid_constant__ ImgPtrs depth, int depth_stride,
    const __grid_constant__ FramePrep prep)
{
    const rgbdfe_frame F = frames.f[blockIdx.x];
    const int n = *F.n;
    const void* img = depth.p[blockIdx.x];
    for (int base = 0; base < n; base += UNPROJ_THREADS * UNPROJ_ILP) {
        float z[UNPROJ_ILP];
#pragma unroll
        for (int u = 0; u < UNPROJ_ILP; u++) {
            const int gi = base + u * UNPROJ_THREADS + threadIdx.x;
            z[u] = 0.f;
            if (gi < n) {
                const size_t di = (size_t)(int)F.kps[gi].y * depth_stride + (int)F.kps[gi].x; // the raw keypoint position
                z[u] = prep.depth16_factor > 0.0f ? (float)static_cast<const unsigned short*>(img)[di] : static_cast<const float*>(img)[di];
            }
        }
#pragma unroll
        for (int u = 0; u < UNPROJ_ILP; u++) {
            const int gi = base + u * UNPROJ_THREADS + threadIdx.x;
            if (gi < n) {
                const float zz = prep.depth16_factor > 0.0f ? __fmul_rn(z[u], prep.depth16_factor) : z[u];
                const float2 un = F.pt_un[gi];
                rgbdfe_point3f P = { 0.f, 0.f, 0.f };
                if (zz > 0) {
                    P.x = __fmul_rn(__fmul_rn(__fsub_rn(un.x, prep.cxf), zz), prep.invfx);
                    P.y = __fmul_rn(__fmul_rn(__fsub_rn(un.y, prep.cyf), zz), prep.invfy);
                    P.z = zz;
                }
                F.xyz[gi] = P;
            }
        }
    }
}

// ---------------------------------------------------------------- launchers
static bool g_pattern_uploaded[64] = { false };

int launch_quadtree(rgbdfe_ctx* c, int count)
{
    const int NC = c->cfg.max_quota + 16;
    int SC = 1;
    while (SC < NC)
        SC <<= 1;
    size_t smem = (size_t)SC * 8 + (size_t)NC * (2 * sizeof(QtNode) + 4 * 4 + 4 * 5) + QT_THREADS * 4;
    CUDA_TRY(c, ensure_dyn_smem(c, k_quadtree, smem));
    ProfScope prof(c, KID_QUADTREE);
    k_quadtree<<<dim3(c->cfg.nlevels, count), QT_THREADS, smem, c->stream>>>(c->d_cfg, c->d_cell_count, c->d_cell_list, c->cell_slot, c->d_cand,
        c->d_label, c->cand_slot, c->d_cand_count, c->d_sel, c->sel_slot, c->d_sel_count, NC, SC);
    c->launches++;
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

int launch_describe(rgbdfe_ctx* c, int count, const FrameBatch& frames, bool unproject, const ImgPtrs& depth, int depth_stride, const FramePrep& prep,
    bool depth_on_host)
{
    if (!g_pattern_uploaded[c->device]) {
        CUDA_TRY(c, cudaMemcpyToSymbolAsync(g_pattern, kBriefPattern, 1024, 0, cudaMemcpyHostToDevice, c->stream));
        g_pattern_uploaded[c->device] = true;
    }
    const int warps = c->kp_cap + 1; // one extra warp publishes the count
    ProfScope prof(c, KID_DESCRIBE);
    k_orient_describe<<<dim3((warps + 7) / 8, count), 256, 0, c->stream>>>(c->d_cfg, c->d_pyr, c->pyr_slot, c->d_blur, c->blur_slot, c->d_sel,
        c->sel_slot, c->d_sel_count, frames, depth, depth_stride, (unproject && !depth_on_host) ? 1 : 0, prep);
    c->launches++;
    if (unproject && depth_on_host) {
        k_unproject_host<<<count, UNPROJ_THREADS, 0, c->stream>>>(frames, depth, depth_stride, prep);
        c->launches++;
    }
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
