// Microbenchmark: latency of single-thread / single-warp sequential FP32 chains in shared memory,
// alone and with the rest of a 512-thread CTA parked at a named barrier.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, long long* cyc, int n, int mode)
{
    extern __shared__ float sm[];
    float* g_w = sm;
    float* g_a = sm + 4096;
    const int tid = threadIdx.x;
    for (int i = tid; i < 4096; i += blockDim.x) { g_w[i] = 1.0f / (1 + i); g_a[i] = 0.5f / (1 + i); }
    __syncthreads();
    if (mode == 2 && tid >= 32 && tid < 64) {
        if (tid == 32) {
            const double* d = reinterpret_cast<const double*>(sm + 8192);
            double sum = 0;
            long long t0 = clock64();
            for (int rep = 0; rep < 4; rep++)
                for (int k = 0; k < n; k += 4) { sum += d[k]; sum += d[k + 1]; sum += d[k + 2]; sum += d[k + 3]; }
            long long t1 = clock64();
            out[100] = (float)sum;
            cyc[2] = t1 - t0;
        }
        return;
    }
    if (mode == 2 && tid >= 64) return;
    if (mode == 1 && tid >= 64) { asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x - 32)); return; }
    if (mode == 1 && tid >= 32) return;
    if (tid == 0) {
        long long t0 = clock64();
        float acc = 0.f;
        float4 w4 = *reinterpret_cast<const float4*>(g_w);
        for (int k = 0; k < n; k += 4) {
            const float4 wn = *reinterpret_cast<const float4*>(g_w + k + 4);
            float4 a4;
            acc += w4.x; a4.x = acc; acc += w4.y; a4.y = acc; acc += w4.z; a4.z = acc; acc += w4.w; a4.w = acc;
            *reinterpret_cast<float4*>(g_a + k) = a4;
            w4 = wn;
        }
        long long t1 = clock64();
        out[0] = acc;
        cyc[0] = t1 - t0;
    }
    __syncwarp();
    if (tid < 32) {
        long long t0 = clock64();
        float m1 = 0, m2 = 0, cv = 0;
        const int c = tid % 9;
        const float* gf = g_w + (c % 3) * 8;
        const float* gt = g_w + (c / 3) * 8 + 64;
        float4 a4 = *reinterpret_cast<const float4*>(g_a), x4 = *reinterpret_cast<const float4*>(gf), y4 = *reinterpret_cast<const float4*>(gt);
        for (int k = 0; k < n; k += 4) {
            const float4 an = *reinterpret_cast<const float4*>(g_a + k + 4), xn = *reinterpret_cast<const float4*>(gf + k + 4), yn = *reinterpret_cast<const float4*>(gt + k + 4);
            const float aa[4] = { a4.x, a4.y, a4.z, a4.w }, xx[4] = { x4.x, x4.y, x4.z, x4.w }, yy[4] = { y4.x, y4.y, y4.z, y4.w };
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float alpha = aa[u];
                const float d1 = xx[u] - m1, d2 = yy[u] - m2;
                const float oma = 1.0f - alpha;
                const float ad2 = alpha * d2;
                cv = oma * (cv + ad2 * d1);
                m1 += alpha * d1;
                m2 += ad2;
            }
            a4 = an; x4 = xn; y4 = yn;
        }
        long long t1 = clock64();
        out[1 + tid] = m1 + m2 + cv;
        if (tid == 0) cyc[1] = t1 - t0;
    }
    if (mode == 1 && tid < 32) asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x - 32));
}
int main()
{
    float* out; long long* cyc;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 64);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
    for (int mode = 0; mode < 3; mode++)
        for (int rep = 0; rep < 3; rep++) {
            k<<<1, mode ? 512 : 32, 60000>>>(out, cyc, 800, mode);
            long long h[3];
            cudaMemcpy(h, cyc, 24, cudaMemcpyDeviceToHost);
            printf("mode %d: prefix %.1f cyc/pt, chain %.1f cyc/pt, dsum %.1f cyc/add (%s)\n", mode, h[0] / 800.0, h[1] / 800.0, h[2] / 3200.0, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
