// Microbenchmark: issue rate of the legacy mma.sync m16n8k32 u8 path (IMMA.16832) on sm_100a, per SM.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__global__ void k(int iters, int* out, long long* cyc)
{
    unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    int c[8][4] = {};
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0 + j), "r"(b1));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < 8; j++)
        s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0)
        cyc[blockIdx.x] = t1 - t0;
}
int main()
{
    int* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 148 * 8);
    for (int warps = 1; warps <= 16; warps *= 2) {
        const int iters = 2000;
        k<<<148, warps * 32>>>(iters, out, cyc);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        const double imma = (double)iters * 8 * warps;
        printf("warps/SM=%2d cycles=%lld  IMMA/clk/SM=%.3f  (u8 MAC/clk/SM=%.0f)\n", warps, h[0], imma / h[0], imma / h[0] * 16 * 8 * 32);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
