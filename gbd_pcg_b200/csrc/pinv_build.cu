// pinv_build.cu -- instantiations and launch planning of the preconditioner builder (pinv_build.cuh)
#include "pinv_build.h"

#include <mutex>
#include <vector>

#include "kernel_table.h"
#include "pinv_build.cuh"

namespace gbdpcg {

static constexpr size_t kMaxSmem = 227 * 1024;
static constexpr size_t kTwoPerSm = 112 * 1024;   // two CTAs per SM: one stages / stores while the other computes

struct PFn { const void *fn; int device; };
static std::mutex g_mu;
static std::vector<PFn> g_ready;

static cudaError_t prepare(const void *fn)
{
    int dev;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto &s : g_ready)
        if (s.fn == fn && s.device == dev) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e != cudaSuccess) return e;
    g_ready.push_back({fn, dev});
    return cudaSuccess;
}

template <typename T>
static const void *specialised_fn(uint32_t nx)
{
    switch (nx) {
#define CASE(NX) case NX: return (const void *)pinv_build_kernel<T, NX>;
        GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
        default: return nullptr;
    }
}

template <typename T>
int launch_pinv_build(const T *d_S, T *d_Pinv, uint32_t nx, uint32_t N, uint32_t batch, int kind, cudaStream_t stream,
                      PinvLaunchInfo *info, bool plan_only)
{
    PinvArgs a;
    a.S = d_S; a.P = d_Pinv; a.mat_stride = 3ull * N * nx * nx; a.N = N; a.batch = batch; a.kind = kind;
    a.use_tma = ((uintptr_t)d_S % 16 == 0 && (uintptr_t)d_Pinv % 16 == 0 && (3ull * nx * nx * sizeof(T)) % 16 == 0) ? 1 : 0;
    int sms = 148, dev = 0;
    if (!plan_only) {
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return (int)e;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return (int)e;
    }
    const void *fn = specialised_fn<T>(nx);
    PinvLaunchInfo li = {};
    if (fn) {
        const int threads = 256;
        // largest chunk of block-rows whose image fits: aim at two CTAs per SM, fall back to one
        int chunk = 0;
        for (size_t budget : {kTwoPerSm, kMaxSmem}) {
            int c = (int)N;
            while (c > 1 && pinv_smem_bytes<T>((int)nx, c, threads) > budget) c = (c + 1) / 2;
            if (pinv_smem_bytes<T>((int)nx, c, threads) <= budget) { chunk = c; break; }
        }
        // few systems: cut them finer so that the whole GPU works on them (latency of a single solve's preconditioner);
        // each chunk re-inverts its two halo blocks, so not below 4 block-rows
        if (chunk > 4 && (unsigned long long)batch * ((N + chunk - 1) / chunk) < (unsigned long long)sms) {
            unsigned long long want = ((unsigned long long)batch * N + 2ull * sms - 1) / (2ull * sms);
            int c = (int)(want < 4 ? 4 : want);
            if (c < chunk) chunk = c;
        }
        if (chunk == 0) return -3;
        a.chunk = (uint32_t)chunk;
        a.nchunks = (N + chunk - 1) / chunk;
        li.specialised = 1; li.threads = threads; li.chunk = chunk;
        li.smem = pinv_smem_bytes<T>((int)nx, chunk, threads);
        li.name = sizeof(T) == 8 ? "pinv_build<f64>" : "pinv_build<f32>";
        const unsigned long long items = (unsigned long long)batch * a.nchunks;
        const int per_sm = li.smem <= kTwoPerSm ? 2 : 1;
        li.grid = (int)(items < (unsigned long long)(sms * per_sm) ? items : (unsigned long long)(sms * per_sm));
    } else {
        li.smem = pinv_generic_smem_bytes<T>(nx);
        if (li.smem > kMaxSmem) return -3;
        fn = (const void *)pinv_build_generic_kernel<T>;
        a.chunk = 1; a.nchunks = N;
        li.specialised = 0; li.threads = 128; li.chunk = 1;
        li.name = sizeof(T) == 8 ? "pinv_build_generic<f64>" : "pinv_build_generic<f32>";
        const unsigned long long items = (unsigned long long)batch * N;
        const unsigned long long cap = (unsigned long long)sms * 8;
        li.grid = (int)(items < cap ? items : cap);
    }
    if (info) *info = li;
    if (plan_only || batch == 0) return 0;
    cudaError_t e = prepare(fn);
    if (e != cudaSuccess) return (int)e;
    uint32_t nxv = nx;
    void *args_s[] = {&a};
    void *args_g[] = {&a, &nxv};
    e = cudaLaunchKernel(fn, dim3(li.grid), dim3(li.threads), li.specialised ? args_s : args_g, li.smem, stream);
    return (int)e;
}

template int launch_pinv_build<float>(const float *, float *, uint32_t, uint32_t, uint32_t, int, cudaStream_t, PinvLaunchInfo *, bool);
template int launch_pinv_build<double>(const double *, double *, uint32_t, uint32_t, uint32_t, int, cudaStream_t, PinvLaunchInfo *, bool);

}  // namespace gbdpcg
