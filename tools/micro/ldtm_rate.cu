// Microbenchmark: read bandwidth of tensor memory (tcgen05.ld.32x32b.x32) per SM on sm_100a, for 1 / 2 / 4 / 8 warps
// per CTA (one CTA per SM, 512 columns allocated).  Bytes per clock and SM = warps * iterations * 4 KB / clocks.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(512, 1) k(int iters, unsigned* out, long long* cyc, int x64)
{
    __shared__ unsigned s_tmem;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem;
    const unsigned taddr = tmem + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * 64);
    unsigned acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        unsigned v[32];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
                     "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                     "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                     "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
                     "=r"(v[31])
                     : "r"(taddr + (unsigned)((i & 1) * 32))
                     : "memory");
        if (x64 || (i & 3) == 3)
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc += v[0] ^ v[31];
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0)
        cyc[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
int main()
{
    unsigned* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 512 * 4);
    cudaMalloc(&cyc, 148 * 8);
    for (int x64 = 0; x64 < 2; x64++)
        for (int warps = 1; warps <= 16; warps *= 2) {
            const int iters = 4096;
            k<<<148, warps * 32>>>(iters, out, cyc, x64);
            cudaDeviceSynchronize();
            long long h = 0;
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            printf("wait_each=%d warps=%2d clocks=%lld  B/clk/SM=%.1f  clk per ld.x32 per warp=%.1f\n", x64, warps, h, (double)warps * iters * 4096.0 / h, (double)h / iters);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
