// K7 on the 5th-generation tensor cores: k_knn2_umma — cv::BFMatcher(NORM_HAMMING)::knnMatch(k = 2) of
// Matcher::match (Features/Matcher.cpp:106-139) as hamming(q, t) = |q| + |t| - 2 (q . t), the q . t of a 128-query x
// 128-train tile being ONE accumulator of tcgen05.mma.kind::i8 (u8 x u8 -> s32, M = 128, N = 128, K = 32 per
// instruction, 8 instructions per tile) held in tensor memory.
//
//   * operands: the 256-bit descriptors expanded to 256 bytes in shared memory (queries 0 / 1, train 0 / 128),
//     K-major without swizzle ("interleaved" canonical layout: 8 rows x 16 bytes core matrices; chunk kc of row r at
//     kc * 2048 + r * 16, so LBO = 2048 (next 16 bytes of K), SBO = 128 (next 8 rows)).  The query tile is expanded
//     once per CTA, the train tiles are double-buffered, their descriptors loaded one tile ahead.
//   * pipeline inside the CTA (no warp specialisation; two CTAs per SM fill each other's gaps): while the tensor
//     pipe multiplies tile i + 1 into TMEM buffer (i + 1) & 1, all 8 warps drain buffer i & 1 with tcgen05.ld
//     (thread = one query row x 64 columns) into a packed top-2 of 16-bit candidate words (distance << 6 | column,
//     two columns per register, VIMNMX.U16x2), widen the tile's four survivors to distance << 16 | train index
//     (unsigned order = OpenCV's (distance, trainIdx) order) and keep the running top-2.
//   * completion of a tile's MMAs is signalled by tcgen05.commit on an mbarrier; waits trap instead of hanging.
// Output format and train-range splits are those of k_knn2_hamming (match.cu), which stays selectable
// (option "knn_umma" = 0) for A/B measurements.  Measured variants that were slower and dropped: per-warp mbarrier
// hand-offs instead of the CTA barrier, 512-thread CTAs (UK_THREADS 512 still builds: the code is generic in UK_PARTS).
#include "match_filter.cuh"
#include "tma.cuh"

namespace rgbdfe {

#define UK_THREADS 256
#define UK_PARTS (UK_THREADS / 128) // threads per query row in the drain (column parts) and per descriptor in the expansion
#define UK_WPP (8 / UK_PARTS)       // 32-bit descriptor words per expansion thread
#define UK_RPW (32 / UK_PARTS)      // descriptor rows per warp in the expansion
#define UK_COLS (UK_N / UK_PARTS)   // accumulator columns per thread and tile
#define UK_M 128          // queries per CTA = TMEM lanes
#define UK_N 128          // train descriptors per tile = TMEM columns per buffer
#define UK_OPERAND_BYTES (UK_M * 256)
#define UK_LBO 2048u      // bytes between the two 16-byte K chunks of one instruction
#define UK_SBO 128u       // bytes between 8-row groups
#define UK_KEY_BIAS 256
#define UK_EMPTY 0x7fffffffu
#define UK_PAD 0x02010000u // words at or above this are not candidates (a real distance + 256 is <= 512)
#define UK_NONE 0x7fffffff
#define UK_SMEM (3 * UK_OPERAND_BYTES + 1024) // query tile + 2 train tiles + alignment slack

// shared-memory matrix descriptor: start address, LBO, SBO in 16-byte units, version 1 (sm_100), no swizzle
__device__ __forceinline__ uint64_t uk_desc(unsigned saddr)
{
    return (uint64_t)((saddr >> 4) & 0x3fffu) | (uint64_t)(UK_LBO >> 4) << 16 | (uint64_t)(UK_SBO >> 4) << 32 | (uint64_t)1 << 46;
}
// instruction descriptor: D = s32 (2 << 4), A = B = u8 (0), both K-major (0), N >> 3 at bit 17, M >> 4 at bit 24
#define UK_IDESC ((2u << 4) | ((unsigned)(UK_N >> 3) << 17) | ((unsigned)(UK_M >> 4) << 24))

__device__ __forceinline__ void uk_mma(unsigned tmem_d, uint64_t da, uint64_t db, unsigned accumulate)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "setp.ne.b32 p, %4, 0;\n"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
                 "}\n" ::"r"(tmem_d),
                 "l"(da), "l"(db), "r"(UK_IDESC), "r"(accumulate)
                 : "memory");
}
// one lane of a converged warp
__device__ __forceinline__ bool uk_elect()
{
    unsigned p;
    asm volatile("{\n"
                 ".reg .pred q;\n"
                 "elect.sync _|q, 0xffffffff;\n"
                 "selp.u32 %0, 1, 0, q;\n"
                 "}\n"
                 : "=r"(p));
    return p != 0;
}
__device__ __forceinline__ void uk_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// a wait that fails loudly instead of hanging the GPU if a completion never arrives
__device__ __forceinline__ void uk_wait(uint64_t* bar, unsigned phase)
{
    unsigned done = 0;
    for (unsigned spin = 0; !done; spin++) {
        asm volatile("{\n"
                     ".reg .pred p;\n"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                     "selp.u32 %0, 1, 0, p;\n"
                     "}\n"
                     : "=r"(done)
                     : "r"(smem_u32(bar)), "r"(phase)
                     : "memory");
        if (spin > (1u << 22))
            __trap();
    }
}
// 32 consecutive TMEM columns of this thread's lane
__device__ __forceinline__ void uk_tmem_ld32(unsigned taddr, unsigned (&v)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
                 "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                 "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                 "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
                 "=r"(v[31])
                 : "r"(taddr)
                 : "memory");
}

// two 16-bit candidates at once (per-lane unsigned min / max)
__device__ __forceinline__ void uk_push2(unsigned p, unsigned& m0, unsigned& m1)
{
    m1 = __vminu2(m1, __vmaxu2(m0, p));
    m0 = __vminu2(m0, p);
}
// 16-bit candidate (distance << 6 | column & 63) -> distance << 16 | train index; 0xffff (no candidate) lands above
// every real word (distance 1023 > 512)
__device__ __forceinline__ unsigned uk_widen(unsigned k, unsigned colbase) { return (k >> 6) << 16 | (colbase + (k & 63u)); }
__device__ __forceinline__ void uk_push(unsigned p, unsigned& m0, unsigned& m1)
{
    m1 = min(m1, max(m0, p));
    m0 = min(m0, p);
}

// Expand 128 descriptors (rows row0 .. row0 + 127 of `desc`, zero beyond `n_valid`) into an operand tile.  A warp
// takes UK_RPW consecutive rows, UK_PARTS threads per row (lane = part * UK_RPW + row): the global loads of a warp are
// one contiguous run, the 8 lanes of a shared-memory store phase write 8 consecutive rows of one 16-byte chunk
// (conflict-free), and the row's population count is a butterfly over the parts.  The order of the 256 positions
// along K is free as long as both operands use the same one: output word j (0..7) of a 32-bit descriptor word w holds
// its bits j, j + 8, j + 16, j + 24, one per byte, i.e. one shift and one mask per 4 bits.  The query tile holds
// 0 / 1; the train tiles (TRAIN) hold 0 / 128 — shift left and mask 0x80808080, the shift being a multiplication that
// ptxas can place on the FMA pipe — so that the accumulator is 128 (q . t), exactly what the 16-bit candidate word
// below subtracts.  TRAIN also leaves the column's base half-word (|t| + 256) << 6 | (column & 63) in s_base (0xffff
// beyond the end of the train set).  The load is a separate step so that the kernel can issue it one tile ahead.
struct UkRow {
    unsigned w[UK_WPP]; // this thread's words of the row's descriptor, zero beyond the end of the set
    bool valid;
};
__device__ __forceinline__ UkRow uk_load(const uint8_t* __restrict__ desc, int row0, int n_valid, int tid)
{
    UkRow R;
    const int lane = tid & 31, r = (tid >> 5) * UK_RPW + (lane & (UK_RPW - 1)), part = lane / UK_RPW;
#pragma unroll
    for (int i = 0; i < UK_WPP; i++)
        R.w[i] = 0u;
    R.valid = row0 + r < n_valid;
    if (R.valid) {
        const uint8_t* p = desc + (size_t)(row0 + r) * 32 + part * (4 * UK_WPP);
        if (UK_WPP == 2) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
            R.w[0] = v.x, R.w[1] = v.y;
        } else {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
            R.w[0] = v.x, R.w[1] = v.y, R.w[UK_WPP - 2] = v.z, R.w[UK_WPP - 1] = v.w;
        }
    }
    return R;
}
template <bool TRAIN>
__device__ __forceinline__ void uk_expand(uint8_t* tile, const UkRow& R, unsigned short* s_base, int tid)
{
    const int lane = tid & 31, r = (tid >> 5) * UK_RPW + (lane & (UK_RPW - 1)), part = lane / UK_RPW;
    uint8_t* dst = tile + (size_t)(2 * UK_WPP * part) * UK_LBO + r * 16;
#pragma unroll
    for (int c = 0; c < 2 * UK_WPP; c++) { // chunk 2 UK_WPP part + c = output words 4 (c & 1) .. + 3 of this thread's word c >> 1
        const unsigned v = R.w[c >> 1];
        const int j0 = 4 * (c & 1);
        uint4 o;
        if (TRAIN)
            o = make_uint4((v << (7 - j0)) & 0x80808080u, (v << (6 - j0)) & 0x80808080u, (v << (5 - j0)) & 0x80808080u, (v << (4 - j0)) & 0x80808080u);
        else
            o = make_uint4((v >> j0) & 0x01010101u, (v >> (j0 + 1)) & 0x01010101u, (v >> (j0 + 2)) & 0x01010101u, (v >> (j0 + 3)) & 0x01010101u);
        *reinterpret_cast<uint4*>(dst + (size_t)c * UK_LBO) = o;
    }
    if (TRAIN) {
        int pop = 0;
#pragma unroll
        for (int i = 0; i < UK_WPP; i++)
            pop += __popc(R.w[i]);
#pragma unroll
        for (int x = UK_RPW; x < 32; x <<= 1)
            pop += __shfl_xor_sync(0xffffffffu, pop, x);
        if (part == 0)
            s_base[r] = R.valid ? (unsigned short)((pop + UK_KEY_BIAS) << 6 | (r & 63)) : (unsigned short)0xffffu;
    }
}

// partial: [pair][split][q] -> int4 (d0, i0, d1, i1), as k_knn2_hamming
__global__ void __launch_bounds__(UK_THREADS, 2) k_knn2_umma(const MatchPair* __restrict__ pairs, int slot0, int4* __restrict__ partial, int q_cap,
    int splits)
{
    const int pair = slot0 + blockIdx.z;
    const MatchPair P = pairs[pair];
    const int nq = *P.nq, nt = *P.nt;
    const int q_base = blockIdx.x * UK_M;
    if (q_base >= nq)
        return;
    extern __shared__ uint8_t uk_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(uk_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB[2] = { smem + UK_OPERAND_BYTES, smem + 2 * UK_OPERAND_BYTES };
    __shared__ __align__(16) unsigned short s_base[3][UK_N]; // tile i uses row i % 3: a fast thread may expand tile i + 2 while others still drain tile i
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ unsigned s_tmem;
    __shared__ uint2 s_merge[UK_PARTS - 1][UK_M];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // train range of this split, in whole tiles
    const int tiles = (nt + UK_N - 1) / UK_N;
    const int tps = (tiles + splits - 1) / splits;
    const int t_begin = min(blockIdx.y * tps * UK_N, nt), t_end = min((blockIdx.y + 1) * tps * UK_N, nt);
    const int T = (t_end - t_begin + UK_N - 1) / UK_N;

    if (warp == 0) { // 2 x 128 accumulator columns; two CTAs per SM use all 512
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
    }
    {
        const UkRow qa = uk_load(P.q_desc, q_base, nq, tid), t0 = uk_load(P.t_desc, t_begin, t_end, tid); // both in flight
        uk_expand<false>(sA, qa, nullptr, tid);
        if (T > 0)
            uk_expand<true>(sB[0], t0, s_base[0], tid);
    }
    UkRow nxt = uk_load(P.t_desc, t_begin + UK_N, t_end, tid); // tile 1 (all zero if there is none)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem;
    const uint64_t dA = uk_desc(smem_u32(sA));
    if (warp == 0 && T > 0 && uk_elect()) { // warp-uniform branch + elect.sync: the descriptors stay in uniform registers
        const uint64_t dB = uk_desc(smem_u32(sB[0]));
#pragma unroll
        for (int s = 0; s < 8; s++) // k-step s = chunks 2s, 2s + 1: start address + 2 * LBO (in 16-byte units)
            uk_mma(tmem, dA + (uint64_t)(s * 2 * (UK_LBO >> 4)), dB + (uint64_t)(s * 2 * (UK_LBO >> 4)), s > 0);
        uk_commit(&s_bar[0]);
    }

    unsigned m0 = UK_EMPTY, m1 = UK_EMPTY;
    const int quad = warp & 3, part = warp >> 2; // TMEM lane quadrant of this warp, column part it drains
    for (int i = 0, sb = 0; i < T; i++, sb = sb == 2 ? 0 : sb + 1) {
        const int b = i & 1, sb1 = sb == 2 ? 0 : sb + 1;
        if (i + 1 < T) {
            // buffer b ^ 1 (smem and TMEM) was last used by tile i - 1, whose MMAs and epilogue are complete
            uk_expand<true>(sB[b ^ 1], nxt, s_base[sb1], tid);
            nxt = uk_load(P.t_desc, t_begin + (i + 2) * UK_N, t_end, tid); // lands while this thread drains tile i
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (warp == 0 && i + 1 < T && uk_elect()) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t dB = uk_desc(smem_u32(sB[b ^ 1]));
#pragma unroll
            for (int s = 0; s < 8; s++)
                uk_mma(tmem + (unsigned)((b ^ 1) * UK_N), dA + (uint64_t)(s * 2 * (UK_LBO >> 4)), dB + (uint64_t)(s * 2 * (UK_LBO >> 4)), s > 0);
            uk_commit(&s_bar[b ^ 1]);
        }
        uk_wait(&s_bar[b], (unsigned)(i >> 1) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // drain: this thread's row (lane quad * 32 + lane), columns [UK_COLS * part, UK_COLS * part + UK_COLS) of buffer b
        const unsigned taddr = tmem + ((unsigned)(quad * 32) << 16) + (unsigned)(b * UK_N + part * UK_COLS);
        unsigned v[UK_COLS / 32][32];
#pragma unroll
        for (int l = 0; l < UK_COLS / 32; l++)
            uk_tmem_ld32(taddr + 32u * l, v[l]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        // Candidates of this tile as 16-bit words, two columns per register: (|t| + 256 - 2 q.t) << 6 | column & 63
        // = base half-word - accumulator (128 q.t <= 32768: no borrow between the halves); unsigned order of a half-word
        // = (distance, column) order.  Packed top-2 with the two-lane min / max instructions: 3 per column pair, on
        // two independent chains.
        const uint4* bp = reinterpret_cast<const uint4*>(&s_base[sb][part * UK_COLS]);
        unsigned M0 = 0xffffffffu, M1 = 0xffffffffu, N0 = 0xffffffffu, N1 = 0xffffffffu;
#pragma unroll
        for (int l = 0; l < UK_COLS / 32; l++)
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint4 bw = bp[4 * l + j];
                uk_push2(bw.x - (v[l][8 * j + 1] * 65536u + v[l][8 * j]), M0, M1);
                uk_push2(bw.y - (v[l][8 * j + 3] * 65536u + v[l][8 * j + 2]), N0, N1);
                uk_push2(bw.z - (v[l][8 * j + 5] * 65536u + v[l][8 * j + 4]), M0, M1);
                uk_push2(bw.w - (v[l][8 * j + 7] * 65536u + v[l][8 * j + 6]), N0, N1);
            }
        {
            const unsigned lo = __vminu2(M0, N0);
            M1 = __vminu2(__vmaxu2(M0, N0), __vminu2(M1, N1));
            M0 = lo;
        }
        // the tile's four survivors as 32-bit words distance << 16 | train index, into the running top-2
        const unsigned colbase = (unsigned)(t_begin + i * UK_N + ((part * UK_COLS) & ~63));
        uk_push(uk_widen(M0 & 0xffffu, colbase), m0, m1);
        uk_push(uk_widen(M0 >> 16, colbase), m0, m1);
        uk_push(uk_widen(M1 & 0xffffu, colbase), m0, m1);
        uk_push(uk_widen(M1 >> 16, colbase), m0, m1);
    }
    // the column parts of a row meet in shared memory; warps 0-3 own the result
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (part > 0)
        s_merge[part - 1][quad * 32 + lane] = make_uint2(m0, m1);
    __syncthreads();
    if (part == 0) {
#pragma unroll
        for (int k = 0; k < UK_PARTS - 1; k++) {
            const uint2 o = s_merge[k][quad * 32 + lane];
            const unsigned lo = min(m0, o.x);
            m1 = min(max(m0, o.x), min(m1, o.y));
            m0 = lo;
        }
        const int q = q_base + quad * 32 + lane;
        if (q < nq) {
            const uint4* p = reinterpret_cast<const uint4*>(P.q_desc + (size_t)q * 32);
            const uint4 w0 = __ldg(p), w1 = __ldg(p + 1);
            const int qpop = __popc(w0.x) + __popc(w0.y) + __popc(w0.z) + __popc(w0.w) + __popc(w1.x) + __popc(w1.y) + __popc(w1.z) + __popc(w1.w);
            int4 r = make_int4(UK_NONE, -1, UK_NONE, -1);
            if (m0 < UK_PAD)
                r.x = qpop + (int)(m0 >> 16) - UK_KEY_BIAS, r.y = (int)(m0 & 0xffffu);
            if (m1 < UK_PAD)
                r.z = qpop + (int)(m1 >> 16) - UK_KEY_BIAS, r.w = (int)(m1 & 0xffffu);
            partial[((size_t)pair * splits + blockIdx.y) * q_cap + q] = r;
        }
    }
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
    }
}

// K7 on tcgen05 for pair slots [slot0, slot0 + npairs); same contract as launch_knn's mma.sync kernel
int launch_knn_umma(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits)
{
    if (!c->knn_umma_ready) { // per context: the attribute belongs to the device the context lives on
        CUDA_TRY(c, cudaFuncSetAttribute(k_knn2_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, UK_SMEM));
        c->knn_umma_ready = true;
    }
    dim3 grd((max_nq + UK_M - 1) / UK_M, splits, npairs);
    k_knn2_umma<<<grd, UK_THREADS, UK_SMEM, c->stream>>>(d_pairs, slot0, c->d_partial, c->kp_cap, splits);
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
