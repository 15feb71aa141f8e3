// Microbenchmark: measured per-SM issue rates of the pipes the path's non-tensor kernels live on (sm_100a), for the
// roofline denominators BASELINE.md section 2 lists: FFMA (FP32 pipe, RANSAC fit), DFMA (FP64 pipe, RANSAC scoring),
// POPC (the integer-pipe Hamming formulation), VIMNMX3 u16x2 (FAST arc min/max), IDP4A (blur taps), LOP3.
// Every thread runs 8 independent dependency chains; 148 CTAs x 1024 threads; ops / clk / SM from clock64 and the
// device-wide rate from CUDA events.  Prints one JSON object.
#include <cuda_runtime.h>

#include <cstdio>

#define ITERS 4096
#define CHAINS 8

template <int OP>
__global__ void __launch_bounds__(1024) k_rate(int iters, unsigned seed, unsigned* out, long long* cyc)
{
    const unsigned t = threadIdx.x + seed;
    float f[CHAINS];
    double d[CHAINS];
    unsigned u[CHAINS];
#pragma unroll
    for (int j = 0; j < CHAINS; j++) {
        f[j] = 1.0f + (float)(t + j) * 1e-6f;
        d[j] = 1.0 + (double)(t + j) * 1e-9;
        u[j] = t * 2654435761u + j;
    }
    const float fa = 1.0000001f, fb = 1e-9f;
    const double da = 1.0000000001, db = 1e-12;
    const unsigned ub = t | 1u;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < CHAINS; j++) {
            if (OP == 0)
                f[j] = fmaf(f[j], fa, fb);
            else if (OP == 1)
                d[j] = fma(d[j], da, db);
            else if (OP == 2)
                u[j] = __popc(u[j]) + ub; // POPC feeding an add: the add goes to the other integer pipe half
            else if (OP == 4)
                u[j] = __dp4a(u[j], ub, u[j]);
            else if (OP == 5)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[j]) : "r"(ub), "r"(seed));
            else if (OP == 6)
                u[j] = __vminu2(__vmaxu2(u[j], ub), seed); // packed u16x2 min / max (lowers to VIMNMX / VIMNMX3 .U16x2)
            else if (OP == 7)
                u[j] = __popc(u[j] ^ ub); // the XOR + POPC pair of the integer Hamming kernel
        }
    }
    const long long t1 = clock64();
    unsigned acc = 0;
#pragma unroll
    for (int j = 0; j < CHAINS; j++)
        acc += u[j] + (unsigned)f[j] + (unsigned)d[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0)
        cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
static void run(const char* name, int sms, unsigned* out, long long* cyc, bool last)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k_rate<OP><<<sms, 1024>>>(64, 1, out, cyc); // warm-up
    cudaEventRecord(a);
    k_rate<OP><<<sms, 1024>>>(ITERS, 1, out, cyc);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    long long h = 0;
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    const double ops_sm = (double)ITERS * CHAINS * 1024;
    printf("  \"%s\": {\"thread_ops_per_clk_per_sm\": %.1f, \"device_Gops_per_s\": %.1f}%s\n", name, ops_sm / (double)h, ops_sm * sms / (ms * 1e-3) / 1e9,
        last ? "" : ",");
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    unsigned* out;
    long long* cyc;
    cudaMalloc(&out, (size_t)sms * 1024 * 4);
    cudaMalloc(&cyc, (size_t)sms * 8);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    printf("{\n  \"device\": \"%s\", \"sms\": %d, \"max_sm_mhz\": %d,\n", p.name, sms, khz / 1000);
    run<0>("ffma_f32", sms, out, cyc, false);
    run<1>("dfma_f64", sms, out, cyc, false);
    run<2>("popc_plus_iadd", sms, out, cyc, false);
    run<7>("xor_popc", sms, out, cyc, false);
    run<6>("vimnmx_u16x2_pair", sms, out, cyc, false);
    run<4>("idp4a", sms, out, cyc, false);
    run<5>("lop3", sms, out, cyc, true);
    printf("}\n");
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        fprintf(stderr, "%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
