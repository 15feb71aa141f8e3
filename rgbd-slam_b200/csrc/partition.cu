// SM partitioning of one B200 between the two halves of the tracking path.
//
// Extraction and kNN are throughput work (grids of thousands of CTAs, every SM busy); the
// filter -> std::sort -> RANSAC chain of a sequence is latency work (1-4 CTAs per kernel, each
// needing ~90 KB of shared memory and 512 threads, ordered pair after pair by the rand() stream
// and the Frame outlier flags).  Sharing the SMs through stream priorities alone doubles the
// chain's latency: a chain CTA becomes runnable only once one SM has drained enough low-priority
// CTAs to fit it.  CUDA green contexts split the SMs instead: a latency partition that the
// chains own, and a throughput partition with the rest; both partitions share the device's
// memory and modules, so frames and workspaces are visible to either.
//
// The driver entry points are resolved through cudaGetDriverEntryPoint: the library does not
// link libcuda (it has to load, and export its symbols, on hosts without a driver).
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace {

typedef CUresult (*fn_cuDeviceGet)(CUdevice*, int);
typedef CUresult (*fn_cuDeviceGetDevResource)(CUdevice, CUdevResource*, CUdevResourceType);
typedef CUresult (*fn_cuDevSmResourceSplitByCount)(CUdevResource*, unsigned int*, const CUdevResource*, CUdevResource*, unsigned int,
    unsigned int);
typedef CUresult (*fn_cuDevResourceGenerateDesc)(CUdevResourceDesc*, CUdevResource*, unsigned int);
typedef CUresult (*fn_cuGreenCtxCreate)(CUgreenCtx*, CUdevResourceDesc, CUdevice, unsigned int);
typedef CUresult (*fn_cuGreenCtxStreamCreate)(CUstream*, CUgreenCtx, unsigned int, int);

struct Driver {
    fn_cuDeviceGet DeviceGet = nullptr;
    fn_cuDeviceGetDevResource DeviceGetDevResource = nullptr;
    fn_cuDevSmResourceSplitByCount DevSmResourceSplitByCount = nullptr;
    fn_cuDevResourceGenerateDesc DevResourceGenerateDesc = nullptr;
    fn_cuGreenCtxCreate GreenCtxCreate = nullptr;
    fn_cuGreenCtxStreamCreate GreenCtxStreamCreate = nullptr;
    bool ok = false;
};

template <class F>
bool entry(const char* name, F& fn)
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p)
        return false;
    fn = reinterpret_cast<F>(p);
    return true;
}

const Driver& driver()
{
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        d.ok = entry("cuDeviceGet", d.DeviceGet) && entry("cuDeviceGetDevResource", d.DeviceGetDevResource)
            && entry("cuDevSmResourceSplitByCount", d.DevSmResourceSplitByCount) && entry("cuDevResourceGenerateDesc", d.DevResourceGenerateDesc)
            && entry("cuGreenCtxCreate", d.GreenCtxCreate) && entry("cuGreenCtxStreamCreate", d.GreenCtxStreamCreate);
    });
    return d;
}

#define MAX_DEVICES 16
struct Partition {
    bool made = false;
    CUgreenCtx latency = nullptr, throughput = nullptr; // live until the process ends
    int latency_sms = 0, throughput_sms = 0;
};
Partition g_part[MAX_DEVICES];
std::mutex g_part_mu;

} // namespace

int rgbdfe_device_partition(int device, int latency_sms, int* latency_got, int* throughput_got)
{
    if (device < 0 || device >= MAX_DEVICES || latency_sms < 1)
        return RGBDFE_E_ARG;
    std::lock_guard<std::mutex> lk(g_part_mu);
    Partition& P = g_part[device];
    if (!P.made) {
        if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess)
            return RGBDFE_E_CUDA;
        const Driver& D = driver();
        if (!D.ok)
            return RGBDFE_E_STATE; // driver without green contexts
        CUdevice dev;
        CUdevResource all, lat, rest;
        unsigned int groups = 1;
        CUdevResourceDesc d_lat = nullptr, d_rest = nullptr;
        if (D.DeviceGet(&dev, device) != CUDA_SUCCESS || D.DeviceGetDevResource(dev, &all, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS)
            return RGBDFE_E_CUDA;
        if ((unsigned)latency_sms >= all.sm.smCount)
            return RGBDFE_E_ARG;
        if (D.DevSmResourceSplitByCount(&lat, &groups, &all, &rest, 0, (unsigned)latency_sms) != CUDA_SUCCESS || groups != 1
            || rest.sm.smCount == 0)
            return RGBDFE_E_CUDA;
        if (D.DevResourceGenerateDesc(&d_lat, &lat, 1) != CUDA_SUCCESS || D.DevResourceGenerateDesc(&d_rest, &rest, 1) != CUDA_SUCCESS)
            return RGBDFE_E_CUDA;
        CUgreenCtx g_lat = nullptr, g_rest = nullptr;
        if (D.GreenCtxCreate(&g_lat, d_lat, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS
            || D.GreenCtxCreate(&g_rest, d_rest, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS)
            return RGBDFE_E_CUDA;
        P.latency = g_lat;
        P.throughput = g_rest;
        P.latency_sms = (int)lat.sm.smCount;
        P.throughput_sms = (int)rest.sm.smCount;
        P.made = true;
    }
    if (latency_got)
        *latency_got = P.latency_sms;
    if (throughput_got)
        *throughput_got = P.throughput_sms;
    return RGBDFE_OK;
}

int rgbdfe_bind_partition(rgbdfe_ctx* c, int role)
{
    if (!c || role < RGBDFE_PART_NONE || role > RGBDFE_PART_LATENCY)
        return RGBDFE_E_ARG;
    cudaSetDevice(c->device);
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    if (c->bulk_stream)
        CUDA_TRY(c, cudaStreamSynchronize(c->bulk_stream));
    cudaStream_t s_main = nullptr, s_bulk = nullptr;
    if (role == RGBDFE_PART_NONE) {
        CUDA_TRY(c, cudaStreamCreateWithFlags(&s_main, cudaStreamNonBlocking));
    } else {
        std::lock_guard<std::mutex> lk(g_part_mu);
        const Partition& P = g_part[c->device < MAX_DEVICES ? c->device : 0];
        if (c->device >= MAX_DEVICES || !P.made) {
            c->err = "rgbdfe_bind_partition: rgbdfe_device_partition has not split this device";
            return RGBDFE_E_STATE;
        }
        const Driver& D = driver();
        int lo = 0, hi = 0;
        CUDA_TRY(c, cudaDeviceGetStreamPriorityRange(&lo, &hi)); // numerically lower = higher priority
        CUstream a = nullptr, b = nullptr;
        const bool latency = role == RGBDFE_PART_LATENCY;
        if (D.GreenCtxStreamCreate(&a, latency ? P.latency : P.throughput, CU_STREAM_NON_BLOCKING, latency ? hi : lo) != CUDA_SUCCESS
            || (latency && D.GreenCtxStreamCreate(&b, P.throughput, CU_STREAM_NON_BLOCKING, lo) != CUDA_SUCCESS)) {
            c->err = "rgbdfe_bind_partition: cuGreenCtxStreamCreate failed";
            return RGBDFE_E_CUDA;
        }
        s_main = (cudaStream_t)a;
        s_bulk = (cudaStream_t)b;
    }
    rgbdfe::release_stream(c);
    c->owns_stream = true;
    if (c->bulk_stream)
        cudaStreamDestroy(c->bulk_stream);
    c->stream = s_main;
    c->bulk_stream = s_bulk;
    return RGBDFE_OK;
}
