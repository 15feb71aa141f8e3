// Probe: 3-D TMA tile load of a u8 tensor (pitch 704) into shared memory, box (BW x BH x 1), arbitrary x offset.
// nvcc -gencode arch=compute_100a,code=sm_100a -I../../rgbd-slam_b200/csrc tma_probe.cu -o tma_probe
#include "tma.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace rgbdfe;
template <int BW, int BH>
__global__ void k_probe(const CUtensorMap* map, int x, int y, int z, uint8_t* out, int mode)
{
    __shared__ __align__(128) uint8_t s[BH][BW];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        if (mode == 1) { // barrier only: a plain arrive completes the phase
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory");
        } else if (mode == 2) { // expect_tx of 0 bytes
            mbar_expect_tx(&bar, 0);
        } else {
            mbar_expect_tx(&bar, BW * BH);
            tma_load_3d(&s[0][0], map, x, y, z, &bar);
        }
    }
    __syncthreads();
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < BW * BH; i += blockDim.x)
        out[i] = (&s[0][0])[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
template <int BW, int BH>
int run(int mode, int x, EncodeTiledFn enc, uint8_t* d_img, const std::vector<uint8_t>& img, int pitch, int rows, int slots, size_t slot_bytes)
{
    CUtensorMap hm;
    const cuuint64_t dims[3] = { (cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)slots };
    const cuuint64_t strides[2] = { (cuuint64_t)pitch, (cuuint64_t)slot_bytes };
    const cuuint32_t box[3] = { BW, BH, 1 }, estr[3] = { 1, 1, 1 };
    CUresult r = enc(&hm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_img, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("box %dx%d encode rc=%d\n", BW, BH, (int)r);
    if (r != CUDA_SUCCESS)
        return 1;
    CUtensorMap* dm;
    uint8_t* d_out;
    cudaMalloc(&dm, sizeof(hm));
    cudaMemcpy(dm, &hm, sizeof(hm), cudaMemcpyHostToDevice);
    cudaMalloc(&d_out, BW * BH);
    const int y = 21, z = 1;
    k_probe<BW, BH><<<1, 128>>>(dm, x, y, z, d_out, mode);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess)
        return 1;
    std::vector<uint8_t> out(BW * BH);
    cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int r2 = 0; r2 < BH; r2++)
        for (int c = 0; c < BW; c++) {
            const int gx = x + c, gy = y + r2;
            const uint8_t want = (gx < pitch && gy < rows) ? img[z * slot_bytes + (size_t)gy * pitch + gx] : 0;
            bad += out[r2 * BW + c] != want;
        }
    printf("  mismatches %ld\n", bad);
    return 0;
}
int main(int argc, char** argv)
{
    const int mode = argc > 1 ? atoi(argv[1]) : 0, x = argc > 2 ? atoi(argv[2]) : 37;
    printf("mode %d x %d\n", mode, x);
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
        printf("no entry point\n");
        return 1;
    }
    EncodeTiledFn enc = (EncodeTiledFn)p;
    const int pitch = 704, rows = 518, slots = 2;
    const size_t slot_bytes = (size_t)pitch * rows + 256 * 7;
    std::vector<uint8_t> img(slot_bytes * slots);
    for (size_t i = 0; i < img.size(); i++)
        img[i] = (uint8_t)(i * 2654435761u >> 13);
    uint8_t* d_img;
    cudaMalloc(&d_img, img.size());
    cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice);
    if (argc > 3 && atoi(argv[3]) == 80)
        run<80, 42>(mode, x, enc, d_img, img, pitch, rows, slots, slot_bytes);
    else
        run<64, 42>(mode, x, enc, d_img, img, pitch, rows, slots, slot_bytes);
    return 0;
}
