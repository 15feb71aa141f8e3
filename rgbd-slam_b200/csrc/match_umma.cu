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
#include <algorithm>

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
// the same with the A operand in tensor memory (lane = row, K bytes packed four per 32-bit column)
__device__ __forceinline__ void uk_mma_ts(unsigned tmem_d, unsigned tmem_a, uint64_t db, unsigned accumulate)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "setp.ne.b32 p, %4, 0;\n"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n"
                 "}\n" ::"r"(tmem_d),
                 "r"(tmem_a), "l"(db), "r"(UK_IDESC), "r"(accumulate)
                 : "memory");
}
// 32 consecutive TMEM columns of this thread's lane <- registers
__device__ __forceinline__ void uk_tmem_st32(unsigned taddr, const unsigned (&v)[32])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
                 "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
                 "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]),
                 "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]),
                 "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
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

// ---------------------------------------------------------------------------------------------------------------
// k_knn2_umma_ws: the same product, warp-specialised and persistent.  What bounds the kernel above is not the tensor
// pipe but the read port of tensor memory: a 128 x 128 s32 accumulator is 64 KB and tcgen05.ld moves 64 B per clock
// and SM, i.e. 1024 clocks per tile against 512 clocks of MMA -- and with every warp doing every job (expand, issue,
// wait, drain) between CTA barriers the port idled about half of the time.  Here one CTA per SM keeps it busy:
//   warps 0-7   DRAIN   thread = one query row x one 64-column half of the tile (two warps per scheduler: one folds
//                       while the other's tcgen05.ld is in flight); tile n: wait acc_full[n % 4], load its 64 columns, free
//                       the accumulator (acc_empty), fold them into the packed top-2.  The halves of a row meet at the item's end.
//   warps 8-15  EXPAND  two groups of 128 threads take alternate tiles (a tile's expansion is latency: 16 stores per thread
//                       and the proxy fence that waits for them); thread = one descriptor row; tile n: wait acc_empty[n % 4]
//                       (the stage's previous tenant has been drained), bits -> bytes into stage n % 4 (descriptor loaded
//                       one tile of the group ahead), arrive b_full.
//   warp  16    ISSUE   one thread: wait b_full, 8 x tcgen05.mma into accumulator n % 4, tcgen05.commit -> acc_full.
// Four operand stages and four accumulators (all 512 TMEM columns) decouple the three; the query tile is double
// buffered across work items (a_full / a_empty), a CTA walks items = (pair, query block, train split) with a grid
// stride, so TMEM is allocated once per CTA and there is no tail of half-empty waves.
// ---------------------------------------------------------------------------------------------------------------
#define WS_THREADS 544 // 8 drain warps + 8 expansion warps + the issue warp
#define WS_ISSUE_WARP 16
#define WS_STAGES 3 // train-tile stages in shared memory = accumulators in tensor memory (3 x 128 columns + 2 x 64 for the query tiles)
#define WS_SMEM (WS_STAGES * UK_OPERAND_BYTES + 1024)

__device__ __forceinline__ void ws_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }

struct WsRow {
    unsigned w[8];
    bool valid;
};
__device__ __forceinline__ WsRow ws_load(const uint8_t* __restrict__ desc, int row, int n_valid)
{
    WsRow R;
    R.valid = row < n_valid;
    uint4 a = make_uint4(0u, 0u, 0u, 0u), b = a;
    if (R.valid) {
        const uint4* p = reinterpret_cast<const uint4*>(desc + (size_t)row * 32);
        a = __ldg(p), b = __ldg(p + 1);
    }
    R.w[0] = a.x, R.w[1] = a.y, R.w[2] = a.z, R.w[3] = a.w, R.w[4] = b.x, R.w[5] = b.y, R.w[6] = b.z, R.w[7] = b.w;
    return R;
}
// One thread expands one descriptor row r of a tile.  K position (chunk c, word jj, byte b) holds bit j + 8 b of descriptor word
// c >> 1, j = 4 (c & 1) + jj, as in uk_expand -- but the two operands carry complementary weights instead of 0 / 1 and 0 / 128:
// a train byte is t_bit * 2^j, which is the descriptor word ANDed with a constant mask and nothing else (no shift: the
// integer ALU pipe is this kernel's bound and the train side is expanded for every tile), a query byte is q_bit * 2^(7 - j)
// (one shift + one mask, once per work item).  Every product is still 128 q_bit t_bit, so the accumulator is 128 (q . t).
template <bool TRAIN>
__device__ __forceinline__ void ws_expand(uint8_t* tile, const WsRow& R, int r, unsigned short* s_base)
{
    uint8_t* dst = tile + r * 16;
#pragma unroll
    for (int c = 0; c < 16; c++) {
        const unsigned v = R.w[c >> 1];
        const int j0 = 4 * (c & 1);
        uint4 o;
        if (TRAIN)
            o = make_uint4(v & (0x01010101u << j0), v & (0x01010101u << (j0 + 1)), v & (0x01010101u << (j0 + 2)), v & (0x01010101u << (j0 + 3)));
        else if (j0 == 0) // bit j -> bit 7 - j: left by 7, 5, 3, 1
            o = make_uint4((v << 7) & 0x80808080u, (v << 5) & 0x40404040u, (v << 3) & 0x20202020u, (v << 1) & 0x10101010u);
        else // j = 4 .. 7: right by 1, 3, 5, 7
            o = make_uint4((v >> 1) & 0x08080808u, (v >> 3) & 0x04040404u, (v >> 5) & 0x02020202u, (v >> 7) & 0x01010101u);
        *reinterpret_cast<uint4*>(dst + (size_t)c * UK_LBO) = o;
    }
    if (TRAIN) {
        int pop = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            pop += __popc(R.w[i]);
        s_base[r] = R.valid ? (unsigned short)((pop + UK_KEY_BIAS) << 6 | (r & 63)) : (unsigned short)0xffffu;
    }
}
// The query row of a work item goes to TENSOR memory (A operand), not to shared memory: a 128 x 128 x 32 multiplication with
// both operands in shared memory reads 8 KB in its 64 clocks, all of the 128 B / clock the SM has, and the expansion stores
// and base-word loads then stretch every multiplication (measured: 138 clocks instead of 64).  TMEM column 8 s + c of the
// row holds K bytes 4 c .. 4 c + 3 of step s, i.e. the words of ws_expand<false> in chunk order.
__device__ __forceinline__ void ws_query_words(const WsRow& R, int half, unsigned (&o)[32])
{
#pragma unroll
    for (int c = 0; c < 8; c++) { // chunks 8 half .. 8 half + 7 = descriptor words 4 half .. 4 half + 3
        const unsigned v = R.w[4 * half + (c >> 1)];
        if ((c & 1) == 0)
            o[4 * c] = (v << 7) & 0x80808080u, o[4 * c + 1] = (v << 5) & 0x40404040u, o[4 * c + 2] = (v << 3) & 0x20202020u, o[4 * c + 3] = (v << 1) & 0x10101010u;
        else
            o[4 * c] = (v >> 1) & 0x08080808u, o[4 * c + 1] = (v >> 3) & 0x04040404u, o[4 * c + 2] = (v >> 5) & 0x02020202u, o[4 * c + 3] = (v >> 7) & 0x01010101u;
    }
}

// 64 accumulator columns of this thread's row against their base half-words -> the running 32-bit top-2.
// The integer ALU pipe (16 lanes per clock and scheduler: every LOP3 / IADD3 / VIMNMX costs a warp two issue clocks) is what
// bounds this kernel, so the fold is written for it: a candidate word is two multiply-adds (FMA pipe), and two candidate
// words enter the packed top-2 together -- sorted against each other first, the new second-best is then one three-input
// minimum: min3(max(m0, lo), m1, hi) -- five VIMNMX per four columns instead of six.
__device__ __forceinline__ unsigned ws_cand(unsigned base2, unsigned v_lo, unsigned v_hi)
{
    unsigned t, p;
    asm("mad.lo.u32 %0, %1, 0xffff0000, %2;" : "=r"(t) : "r"(v_hi), "r"(base2)); // base - (v_hi << 16)
    asm("mad.lo.u32 %0, %1, 0xffffffff, %2;" : "=r"(p) : "r"(v_lo), "r"(t));     //      - v_lo
    return p;
}
__device__ __forceinline__ void ws_push4(unsigned p, unsigned q, unsigned& m0, unsigned& m1)
{
    const unsigned lo = __vminu2(p, q), hi = __vmaxu2(p, q);
    m1 = __vimin3_u16x2(__vmaxu2(m0, lo), m1, hi);
    m0 = __vminu2(m0, lo);
}
__device__ __forceinline__ void ws_fold64(const unsigned (&v)[2][32], const unsigned short* base64, unsigned colbase, unsigned& m0, unsigned& m1)
{
    const uint4* bp = reinterpret_cast<const uint4*>(base64);
    unsigned M0 = 0xffffffffu, M1 = 0xffffffffu, N0 = 0xffffffffu, N1 = 0xffffffffu;
#pragma unroll
    for (int l = 0; l < 2; l++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint4 bw = bp[4 * l + j];
            ws_push4(ws_cand(bw.x, v[l][8 * j], v[l][8 * j + 1]), ws_cand(bw.y, v[l][8 * j + 2], v[l][8 * j + 3]), M0, M1);
            ws_push4(ws_cand(bw.z, v[l][8 * j + 4], v[l][8 * j + 5]), ws_cand(bw.w, v[l][8 * j + 6], v[l][8 * j + 7]), N0, N1);
        }
    const unsigned lo = __vminu2(M0, N0);
    M1 = __vminu2(__vmaxu2(M0, N0), __vminu2(M1, N1));
    M0 = lo;
    uk_push(uk_widen(M0 & 0xffffu, colbase), m0, m1);
    uk_push(uk_widen(M0 >> 16, colbase), m0, m1);
    uk_push(uk_widen(M1 & 0xffffu, colbase), m0, m1);
    uk_push(uk_widen(M1 >> 16, colbase), m0, m1);
}

struct WsItem { // what a role needs to know about a work item, resolved once per CTA (the counts live in device memory)
    const uint8_t* q_desc;
    const uint8_t* t_desc;
    int nq, nt;
};
#define WS_TABLE 256

// ring position of a tile: stage (0 .. WS_STAGES - 1), base-word slot (0 .. 2 WS_STAGES - 1), use count of the stage
struct WsRing {
    int st = 0, sb = 0;
    unsigned use = 0;
    __device__ __forceinline__ void step()
    {
        sb = sb + 1 == 2 * WS_STAGES ? 0 : sb + 1;
        if (++st == WS_STAGES)
            st = 0, use++;
    }
};

__global__ void __launch_bounds__(WS_THREADS, 1) k_knn2_umma_ws(const MatchPair* __restrict__ pairs, int slot0, int npairs, int qblocks, int splits,
    int4* __restrict__ partial, int q_cap, int dbg, long long* __restrict__ tl)
{
    // developer timeline (rgbdfe_debug_timeline): CTA 0 stamps clock64 at the six hand-offs of its first 64 tiles
#define WS_STAMP(ev, cond)                                   \
    do {                                                     \
        if (tl && blockIdx.x == 0 && (cond) && seq < 64)     \
            tl[(ev) * 64 + seq] = clock64();                 \
    } while (0)
    extern __shared__ uint8_t uk_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(uk_smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(16) unsigned short s_base[2 * WS_STAGES][UK_N]; // a tile's base words outlive its accumulator by a fold
    __shared__ __align__(8) uint64_t a_full[2], a_empty[2], b_full[WS_STAGES], acc_full[WS_STAGES], acc_empty[WS_STAGES];
    __shared__ unsigned s_tmem;
    __shared__ uint2 s_merge[2][UK_M];
    __shared__ WsItem s_items[WS_TABLE];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_items = npairs * qblocks * splits;
    if (tid == 0) {
        for (int x = 0; x < 2; x++) {
            mbar_init(&a_full[x], 128);
            mbar_init(&a_empty[x], 1);
        }
        for (int st = 0; st < WS_STAGES; st++) {
            mbar_init(&b_full[st], 128);
            mbar_init(&acc_full[st], 1);
            mbar_init(&acc_empty[st], 256);
        }
    }
    if (warp == WS_ISSUE_WARP) { // 512 columns: three 128-column accumulators + two 64-column query tiles
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the CTA's first WS_TABLE items, resolved by as many threads at once: one latency chain per CTA instead of one per item and role
    for (int k = tid; k < WS_TABLE; k += WS_THREADS) {
        const long long item = (long long)blockIdx.x + (long long)k * gridDim.x;
        if (item < n_items) {
            const MatchPair P = pairs[slot0 + (int)(item / (splits * qblocks))];
            s_items[k] = WsItem { P.q_desc, P.t_desc, *P.nq, *P.nt };
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem;
    const unsigned tmem_a = tmem + (unsigned)(WS_STAGES * UK_N); // query tiles behind the accumulators
    unsigned seq = 0, ait = 0; // tiles / items with tiles handled so far (identical in every role)
    WsRing ring;               // ring position of tile `seq`

    int k_item = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, k_item++) {
        const int split = item % splits, qb = (item / splits) % qblocks, pair = slot0 + item / (splits * qblocks);
        WsItem I;
        if (k_item < WS_TABLE)
            I = s_items[k_item];
        else {
            const MatchPair P = pairs[pair];
            I = WsItem { P.q_desc, P.t_desc, *P.nq, *P.nt };
        }
        const int nq = I.nq, nt = I.nt;
        const int q_base = qb * UK_M;
        if (q_base >= nq)
            continue;
        const int tiles = (nt + UK_N - 1) / UK_N, tps = (tiles + splits - 1) / splits;
        const int t_begin = min(split * tps * UK_N, nt), t_end = min((split + 1) * tps * UK_N, nt);
        const int T = (t_end - t_begin + UK_N - 1) / UK_N;
        const int x = ait & 1;
        if (warp < 8) {
            // ---------------------------------------------------------------- DRAIN
            unsigned m0 = UK_EMPTY, m1 = UK_EMPTY;
            const int quad = warp & 3, half = warp >> 2, row = quad * 32 + lane;
            const int q = q_base + row;
            uint4 w0 = make_uint4(0u, 0u, 0u, 0u), w1 = w0;
            if (half == 0 && q < nq) { // |q| for the result; lands while the first tile is produced
                const uint4* p = reinterpret_cast<const uint4*>(I.q_desc + (size_t)q * 32);
                w0 = __ldg(p), w1 = __ldg(p + 1);
            }
            for (int i = 0; i < T; i++, seq++, ring.step()) {
                uk_wait(&acc_full[ring.st], ring.use & 1u);
                WS_STAMP(4, tid == 0);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned taddr = tmem + ((unsigned)(quad * 32) << 16) + (unsigned)(ring.st * UK_N + half * 64);
                unsigned v[2][32];
                uk_tmem_ld32(taddr, v[0]);
                uk_tmem_ld32(taddr + 32u, v[1]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                ws_arrive(&acc_empty[ring.st]); // the accumulator (and its operand stage) may be refilled
                if (!(dbg & 1))
                    ws_fold64(v, &s_base[ring.sb][half * 64], (unsigned)(t_begin + i * UK_N + half * 64), m0, m1);
                else
                    m0 += v[0][0] + v[1][31];
                WS_STAMP(5, tid == 0);
            }
            if (T > 0) { // the two column halves of a row meet in shared memory (double buffered by item parity)
                if (half == 1)
                    s_merge[x][row] = make_uint2(m0, m1);
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (half == 0) {
                    const uint2 o = s_merge[x][row];
                    const unsigned lo = min(m0, o.x);
                    m1 = min(max(m0, o.x), min(m1, o.y));
                    m0 = lo;
                }
            }
            if (half == 0 && q < nq) {
                const int qpop = __popc(w0.x) + __popc(w0.y) + __popc(w0.z) + __popc(w0.w) + __popc(w1.x) + __popc(w1.y) + __popc(w1.z) + __popc(w1.w);
                int4 r = make_int4(UK_NONE, -1, UK_NONE, -1);
                if (m0 < UK_PAD)
                    r.x = qpop + (int)(m0 >> 16) - UK_KEY_BIAS, r.y = (int)(m0 & 0xffffu);
                if (m1 < UK_PAD)
                    r.z = qpop + (int)(m1 >> 16) - UK_KEY_BIAS, r.w = (int)(m1 & 0xffffu);
                partial[((size_t)pair * splits + split) * q_cap + q] = r;
            }
        } else if (warp < 16) {
            // ---------------------------------------------------------------- EXPAND: two groups of 128 threads take alternate tiles
            const int g = (warp - 8) >> 2, e = tid - 256 - 128 * g;
            if (T > 0) {
                if ((int)(ait & 1u) == g) { // the item's query tile -> tensor memory (this warp's lane quadrant is warp % 4 = e / 32)
                    const WsRow qa = ws_load(I.q_desc, q_base + e, nq);
                    uk_wait(&a_empty[x], ((ait >> 1) & 1u) ^ 1u); // the query buffer's previous item has finished multiplying
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const unsigned ta = tmem_a + ((unsigned)((e >> 5) * 32) << 16) + (unsigned)(x * 64);
                    unsigned o[32];
                    ws_query_words(qa, 0, o);
                    uk_tmem_st32(ta, o);
                    ws_query_words(qa, 1, o);
                    uk_tmem_st32(ta + 32u, o);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    ws_arrive(&a_full[x]);
                }
                WsRing rg = ring;
                int i = (int)((seq ^ (unsigned)g) & 1u); // this group's first tile of the item: global tile parity == g
                if (i)
                    rg.step();
                WsRow nxt = ws_load(I.t_desc, t_begin + i * UK_N + e, t_end);
                for (; i < T; i += 2) {
                    const unsigned sq = seq + (unsigned)i;
                    const WsRow cur = nxt;
                    nxt = ws_load(I.t_desc, t_begin + (i + 2) * UK_N + e, t_end); // this group's next tile (all zero past the end)
                    uk_wait(&acc_empty[rg.st], (rg.use & 1u) ^ 1u);
                    if (tl && blockIdx.x == 0 && e == 0 && sq < 64)
                        tl[0 * 64 + sq] = clock64();
                    if (!(dbg & 2))
                        ws_expand<true>(smem + rg.st * UK_OPERAND_BYTES, cur, e, s_base[rg.sb]);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    ws_arrive(&b_full[rg.st]);
                    if (tl && blockIdx.x == 0 && e == 0 && sq < 64)
                        tl[1 * 64 + sq] = clock64();
                    rg.step();
                    rg.step();
                }
                seq += (unsigned)T;
                for (int i2 = 0; i2 < T; i2++)
                    ring.step();
            }
        } else if (tid == WS_ISSUE_WARP * 32) {
            // ---------------------------------------------------------------- ISSUE
            if (T > 0) {
                uk_wait(&a_full[x], (ait >> 1) & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned tA = tmem_a + (unsigned)(x * 64);
                for (int i = 0; i < T; i++, seq++, ring.step()) {
                    uk_wait(&b_full[ring.st], ring.use & 1u);
                    WS_STAMP(2, true);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint64_t dB = uk_desc(smem_u32(smem + ring.st * UK_OPERAND_BYTES));
                    if (!(dbg & 4)) {
#pragma unroll
                        for (int k = 0; k < 8; k++) { // k-step k: TMEM columns 8 k .. 8 k + 7 of the query tile, chunks 2 k, 2 k + 1 of the train stage
                            uk_mma_ts(tmem + (unsigned)(ring.st * UK_N), tA + (unsigned)(8 * k), dB + (uint64_t)(k * 2 * (UK_LBO >> 4)), k > 0);
                            if (tl && blockIdx.x == 0 && seq >= 20 && seq < 24)
                                tl[400 + (seq - 20) * 8 + k] = clock64();
                        }
                    }
                    uk_commit(&acc_full[ring.st]);
                    WS_STAMP(3, true);
                }
                uk_commit(&a_empty[x]); // every multiplication that reads this query buffer has completed
            }
        } else {
            seq += (unsigned)T; // idle lanes of the issue warp keep the counters in step
            for (int i2 = 0; i2 < T; i2++)
                ring.step();
        }
        if (T > 0)
            ait++;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == WS_ISSUE_WARP) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// K7 on tcgen05 for pair slots [slot0, slot0 + npairs); same contract as launch_knn's mma.sync kernel
int launch_knn_umma(rgbdfe_ctx* c, const MatchPair* d_pairs, int slot0, int npairs, int max_nq, int splits)
{
    if (c->knn_umma == 2) { // the warp-specialised persistent form
        CUDA_TRY(c, ensure_dyn_smem(c, k_knn2_umma_ws, WS_SMEM));
        if (c->sm_count == 0)
            CUDA_TRY(c, cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, c->device));
        const int qblocks = (max_nq + UK_M - 1) / UK_M;
        const int n_items = npairs * qblocks * splits;
        k_knn2_umma_ws<<<std::min(n_items, c->sm_count), WS_THREADS, WS_SMEM, c->stream>>>(d_pairs, slot0, npairs, qblocks, splits, c->d_partial, c->kp_cap, c->knn_dbg, c->dbg_on ? c->d_dbg : nullptr);
        CUDA_TRY(c, cudaGetLastError());
        return RGBDFE_OK;
    }
    CUDA_TRY(c, ensure_dyn_smem(c, k_knn2_umma, UK_SMEM));
    dim3 grd((max_nq + UK_M - 1) / UK_M, splits, npairs);
    k_knn2_umma<<<grd, UK_THREADS, UK_SMEM, c->stream>>>(d_pairs, slot0, c->d_partial, c->kp_cap, splits);
    CUDA_TRY(c, cudaGetLastError());
    return RGBDFE_OK;
}

} // namespace rgbdfe
