// tmem_rate.cu -- can tensor memory (256 KB per SM, idle in this solver) hold one of the two matrices of a system?
// Measures, on one SM, the rate at which warps can stream thread-private data TMEM -> registers with
// tcgen05.ld.32x32b (each lane reads its own TMEM lane: the layout a row-per-thread mat-vec would use), next to the
// shared-memory LDS.128 rate the window loads of the solver already use, and both at once (different warps), to see
// whether the two paths add up or share a port.  Round trip tcgen05.st -> tcgen05.ld is checked first.
//
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tmem_rate tmem_rate.cu      run: ./tmem_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define R16(v) "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}"

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%16], {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15};" ::"r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                 "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct Result {
    long long clocks;
    unsigned long long tmem_bytes, lds_bytes;
    uint32_t errors, checksum;
};

// mode bit 0: warps in `tmem_mask` stream TMEM; mode bit 1: warps in `lds_mask` stream shared memory (LDS.128, per-lane
// addresses in groups of 6 lanes like the solver's windows).  CHUNKS = x16 loads in flight before a wait::ld.
template <int CHUNKS>
__global__ void __launch_bounds__(256, 1) tmem_rate_kernel(int reps, uint32_t tmem_mask, uint32_t lds_mask, Result *out)
{
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float win[2048];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int i = t; i < 2048; i += blockDim.x) win[i] = (float)i;
    if (warp == 0) {
        uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base_s);
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_base_s;
    // warp w owns TMEM lanes 32*(w%4) .. +31 (the hardware's rule) and columns 256*(w/4) .. +255
    const uint32_t my = base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(256 * (warp >> 2));

    // ---- round trip: value = f(thread, column) ----
    uint32_t v[16];
    for (int c = 0; c < 256; c += 16) {
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = (uint32_t)t * 1000u + (uint32_t)(c + i);
        tmem_st16(my + c, v);
    }
    tmem_wait_st();
    uint32_t errors = 0;
    for (int c = 0; c < 256; c += 16) {
        tmem_ld16(my + c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; i++) errors += (v[i] != (uint32_t)t * 1000u + (uint32_t)(c + i));
    }
    __syncthreads();

    const bool do_tmem = (tmem_mask >> warp) & 1u, do_lds = (lds_mask >> warp) & 1u;
    uint32_t acc = 0;
    float facc = 0.f;
    unsigned long long nt = 0, nl = 0;
    const float4 *w4 = reinterpret_cast<const float4 *>(win) + (lane / 6) * 12;
    __syncthreads();
    const long long c0 = clock64();
    for (int r = 0; r < reps; r++) {
        if (do_tmem) {
            // 256 columns = 16 x16-loads, CHUNKS of them in flight per wait
#pragma unroll 1
            for (int c = 0; c < 256; c += 16 * CHUNKS) {
                uint32_t u[CHUNKS][16];
#pragma unroll
                for (int k = 0; k < CHUNKS; k++) tmem_ld16(my + c + 16 * k, u[k]);
                tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < CHUNKS; k++)
#pragma unroll
                    for (int i = 0; i < 16; i++) acc += u[k][i];
            }
            nt += 256ull * 4ull;
        }
        if (do_lds) {
            // the same 1024 bytes per thread from shared memory, 16 bytes per load
#pragma unroll 8
            for (int i = 0; i < 64; i++) {
                const float4 q = w4[(i + r) & 63];
                facc += q.x + q.y + q.z + q.w;
            }
            nl += 1024ull;
        }
    }
    __syncthreads();
    const long long c1 = clock64();
    // totals over the CTA
    __shared__ unsigned long long tot[2];
    __shared__ uint32_t errs, chk;
    if (t == 0) { tot[0] = tot[1] = 0; errs = 0; chk = 0; }
    __syncthreads();
    atomicAdd(&tot[0], nt);
    atomicAdd(&tot[1], nl);
    atomicAdd(&errs, errors);
    atomicAdd(&chk, acc + (uint32_t)facc);
    __syncthreads();
    if (t == 0) {
        out->clocks = c1 - c0;
        out->tmem_bytes = tot[0];
        out->lds_bytes = tot[1];
        out->errors = errs;
        out->checksum = chk;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512) : "memory");
}

template <int CHUNKS>
static void run(const char *name, uint32_t tmem_mask, uint32_t lds_mask)
{
    Result *d, h;
    cudaMalloc(&d, sizeof(Result));
    const int reps = 2000;
    tmem_rate_kernel<CHUNKS><<<1, 256>>>(10, tmem_mask, lds_mask, d);       // warm
    tmem_rate_kernel<CHUNKS><<<1, 256>>>(reps, tmem_mask, lds_mask, d);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-46s chunks/wait %d: %s, round-trip errors %u, %9lld clk, TMEM %7.1f B/clk, LDS %7.1f B/clk\n", name, CHUNKS,
           cudaGetErrorString(e), h.errors, h.clocks, (double)h.tmem_bytes / h.clocks, (double)h.lds_bytes / h.clocks);
    cudaFree(d);
}

int main()
{
    run<1>("TMEM, 1 warp", 0x01, 0);
    run<4>("TMEM, 1 warp", 0x01, 0);
    run<1>("TMEM, 4 warps (one per lane quarter)", 0x0f, 0);
    run<4>("TMEM, 4 warps (one per lane quarter)", 0x0f, 0);
    run<1>("TMEM, 8 warps", 0xff, 0);
    run<2>("TMEM, 8 warps", 0xff, 0);
    run<4>("TMEM, 8 warps", 0xff, 0);
    run<4>("TMEM, 2 warps on the SAME lane quarter", 0x11, 0);
    run<1>("LDS.128, 4 warps", 0, 0x0f);
    run<1>("LDS.128, 8 warps", 0, 0xff);
    run<4>("TMEM warps 0-3 + LDS.128 warps 4-7", 0x0f, 0xf0);
    run<4>("TMEM 8 warps + LDS.128 8 warps (same warps)", 0xff, 0xff);
    return 0;
}
