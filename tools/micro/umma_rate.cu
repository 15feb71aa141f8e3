// Microbenchmark: dispatch rate of tcgen05.mma (cta_group::1, M = 128, N = 128 / 256, both operands in shared memory,
// K = 32 bytes per instruction) for kind::i8, kind::f8f6f4 (e4m3 x e4m3 -> f32) and kind::f16 (bf16, K = 16 elements)
// on sm_100a: one thread per SM issues `n` multiplications back to back into one accumulator, then commits and waits.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(unsigned saddr) { return (uint64_t)((saddr >> 4) & 0x3fffu) | (uint64_t)(2048u >> 4) << 16 | (uint64_t)(128u >> 4) << 32 | (uint64_t)1 << 46; }
template <int KIND>
__device__ __forceinline__ void mma(unsigned d, uint64_t a, uint64_t b, unsigned idesc, unsigned acc)
{
    if (KIND == 0)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else if (KIND == 1)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
template <int KIND>
__global__ void __launch_bounds__(128, 1) k(int n, int N, long long* cyc)
{
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ unsigned s_tmem;
    __shared__ __align__(8) uint64_t bar;
    uint8_t* base = (uint8_t*)(((uintptr_t)sm + 1023) & ~(uintptr_t)1023);
    for (int i = threadIdx.x; i < 3 * 32768 / 4; i += 128)
        ((unsigned*)base)[i] = 0;
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem;
    // instruction descriptor: D format bits 4-5 (1 = f32, 2 = s32), A format bits 7-9, B format bits 10-12, N >> 3 at 17, M >> 4 at 24
    unsigned idesc = ((unsigned)(N >> 3) << 17) | ((unsigned)(128 >> 4) << 24);
    if (KIND == 0)
        idesc |= 2u << 4; // s32, u8 x u8
    else if (KIND == 1)
        idesc |= 1u << 4; // f32, e4m3 x e4m3 (format 0)
    else
        idesc |= (1u << 4) | (1u << 7) | (1u << 10); // f32, bf16 x bf16
    if (threadIdx.x == 0) {
        const uint64_t dA = desc(smem_u32(base)), dB = desc(smem_u32(base + 32768));
        const long long t0 = clock64();
        for (int i = 0; i < n; i++)
            mma<KIND>(tmem, dA + (uint64_t)((i & 7) * 2 * (2048 >> 4)), dB + (uint64_t)((i & 7) * 2 * (2048 >> 4)), idesc, i > 0);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        unsigned done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
        cyc[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
template <int KIND>
static void run(const char* name, int N, long long* cyc)
{
    cudaFuncSetAttribute(k<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768 + 1024);
    const int n = 4096;
    k<KIND><<<148, 128, 3 * 32768 + 1024>>>(64, N, cyc);
    k<KIND><<<148, 128, 3 * 32768 + 1024>>>(n, N, cyc);
    cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)h / n;
    printf("%-22s N=%3d  clocks per instruction %.1f  -> %.0f MAC/clk/SM (K = %d per instruction)  %s\n", name, N, per, 128.0 * N * (KIND == 2 ? 16 : 32) / per,
        KIND == 2 ? 16 : 32, cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    long long* cyc;
    cudaMalloc(&cyc, 148 * 8);
    for (int N = 128; N <= 256; N += 128) {
        run<0>("kind::i8 (u8 x u8)", N, cyc);
        run<1>("kind::f8f6f4 (e4m3)", N, cyc);
        run<2>("kind::f16 (bf16)", N, cyc);
    }
    return 0;
}
