// Micro-benchmark: issue rate of FFMA vs FFMA2 (fma.rn.f32x2) vs DFMA on one SM, 1/2/4 warps per sub-partition,
// 8 independent accumulator chains per thread.  Prints SM-clock cycles per warp-instruction per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float *out, int iters, float a, float b, long long *cyc)
{
    float2 acc[8]; double dacc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f); dacc[i] = threadIdx.x + i; }
    float2 m = make_float2(a, b), x = make_float2(b, a);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if constexpr (MODE == 0) { acc[i].x = fmaf(m.x, x.x, acc[i].x); }
                else if constexpr (MODE == 1) { acc[i] = __ffma2_rn(m, x, acc[i]); }
                else if constexpr (MODE == 2) { dacc[i] = fma((double)a, (double)b, dacc[i]); }
                else { acc[i].x = fmaf(m.x, x.x, acc[i].x); acc[i].y = fmaf(m.y, x.y, acc[i].y); }
            }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y + (float)dacc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main()
{
    float *d; long long *c; cudaMalloc(&d, 1024 * 4); cudaMalloc(&c, 8);
    const int iters = 20000;
    const char *names[] = {"FFMA (scalar)", "FFMA2 (packed)", "DFMA", "2 x FFMA"};
    for (int mode = 0; mode < 4; mode++)
        for (int nt : {128, 256, 512, 1024}) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms = 0;
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                if (mode == 0) k<0><<<1, nt>>>(d, iters, 1.0001f, 0.9999f, c);
                if (mode == 1) k<1><<<1, nt>>>(d, iters, 1.0001f, 0.9999f, c);
                if (mode == 2) k<2><<<1, nt>>>(d, iters, 1.0001f, 0.9999f, c);
                if (mode == 3) k<3><<<1, nt>>>(d, iters, 1.0001f, 0.9999f, c);
                cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
            }
            long long cy; cudaMemcpy(&cy, c, 8, cudaMemcpyDeviceToHost);
            const double ninst = (double)iters * 32 * (mode == 3 ? 2 : 1);   // per warp
            const int wps = nt / 128;
            printf("%-16s %4d threads (%d warps/SMSP): %.2f clk per warp-inst per SMSP (clock64), %.3f ns per warp-inst per SMSP (events) -> %.2f GHz\n",
                   names[mode], nt, wps, (double)cy / (ninst * wps), ms * 1e6 / (ninst * wps), (double)cy / (ms * 1e6));
        }
    return 0;
}
