// Micro-benchmark: FFMA2 issue rate when every instruction reads three DIFFERENT 64-bit register pairs
// (as in the block-row mat-vec: matrix pair, vector pair, accumulator) vs reused operands; also scalar FFMA.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float *out, const float *in, int iters, long long *cyc)
{
    float2 m[32], x[8], acc[4];
#pragma unroll
    for (int i = 0; i < 32; i++) m[i] = make_float2(in[threadIdx.x + i], in[threadIdx.x + 2 * i + 1]);
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = make_float2(in[i + 7], in[i + 3]);
#pragma unroll
    for (int i = 0; i < 4; i++) acc[i] = make_float2(0.f, 0.f);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 32; i++) {
            if constexpr (MODE == 0) acc[i & 3] = __ffma2_rn(m[i], x[i & 7], acc[i & 3]);
            else if constexpr (MODE == 1) acc[i & 3] = __ffma2_rn(m[0], x[0], acc[i & 3]);
            else { acc[i & 3].x = fmaf(m[i].x, x[i & 7].x, acc[i & 3].x); acc[i & 3].y = fmaf(m[i].y, x[i & 7].y, acc[i & 3].y); }
        }
#pragma unroll
        for (int i = 0; i < 8; i++) x[i].x += 1e-9f;   // keep the loop honest
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc[0].x + acc[1].y + acc[2].x + acc[3].y;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main()
{
    float *d, *in; long long *c; cudaMalloc(&d, 1024 * 4); cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0, 4096 * 4); cudaMalloc(&c, 8);
    const int iters = 20000;
    const char *names[] = {"FFMA2 distinct operands", "FFMA2 reused operands", "2xFFMA distinct operands"};
    for (int mode = 0; mode < 3; mode++)
        for (int nt : {128, 256}) {
            for (int rep = 0; rep < 2; rep++) {
                if (mode == 0) k<0><<<1, nt>>>(d, in, iters, c);
                if (mode == 1) k<1><<<1, nt>>>(d, in, iters, c);
                if (mode == 2) k<2><<<1, nt>>>(d, in, iters, c);
                cudaDeviceSynchronize();
            }
            long long cy; cudaMemcpy(&cy, c, 8, cudaMemcpyDeviceToHost);
            printf("%-26s %d warps/SMSP: %.2f clk per 2 FMAs (per warp-instruction-pair) per SMSP\n", names[mode], nt / 128, (double)cy / ((double)iters * 32 * (nt / 128)));
        }
    return 0;
}
