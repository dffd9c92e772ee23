// Micro-benchmark: issue rate of the packed multiply-adds of the four-rows-per-thread mat-vec (pcg_tmem4.cuh) under different
// operand arrangements.  84 matrix pairs live in registers (one mat-vec's worth), the vector quads change every round.
//   MODE 0  column pairs (what the kernel does):   acc[row] = ffma2((m[row][c], m[row][c+1]), (x[c], x[c+1]), acc[row]); x shared by 4 rows
//   MODE 1  the same arithmetic with scalar FFMA
//   MODE 2  row pairs:   acc[rp] = ffma2((m[2rp][c], m[2rp+1][c]), (x[c], x[c]), acc[rp]);  the vector pair is a duplicated scalar
//   MODE 3  column pairs, 2 independent chains per row (8 accumulators)
// Prints clocks per mat-vec (84 FFMA2 or 168 FFMA per thread) per sub-partition for 1 and 2 warps per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float *out, const float *in, int iters, long long *cyc)
{
    float2 m[84];
    float4 x[3];
#pragma unroll
    for (int i = 0; i < 84; i++) m[i] = make_float2(in[threadIdx.x + i], in[threadIdx.x + 2 * i + 1]);
#pragma unroll
    for (int i = 0; i < 3; i++) x[i] = make_float4(in[i + 7], in[i + 3], in[i + 11], in[i + 5]);
    float2 acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = make_float2(0.f, 0.f);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        // 21 column pairs (= 10.5 quads) x 4 rows
#pragma unroll
        for (int q = 0; q < 21; q++) {
            const float4 xq = x[q % 3];
            const float2 xp = (q & 1) ? make_float2(xq.z, xq.w) : make_float2(xq.x, xq.y);
#pragma unroll
            for (int row = 0; row < 4; row++) {
                const float2 mm = m[4 * q + row];
                if constexpr (MODE == 0) acc[row] = __ffma2_rn(mm, xp, acc[row]);
                else if constexpr (MODE == 1) { acc[row].x = fmaf(mm.x, xp.x, acc[row].x); acc[row].y = fmaf(mm.y, xp.y, acc[row].y); }
                else if constexpr (MODE == 2) acc[row] = __ffma2_rn(mm, (row & 1) ? make_float2(xp.y, xp.y) : make_float2(xp.x, xp.x), acc[row]);
                else acc[row + 4 * (q & 1)] = __ffma2_rn(mm, xp, acc[row + 4 * (q & 1)]);
            }
        }
#pragma unroll
        for (int i = 0; i < 3; i++) { x[i].x += 1e-9f; x[i].z += 1e-9f; }   // keep the loop honest
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main()
{
    float *d, *in; long long *c; cudaMalloc(&d, 1024 * 4); cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0, 4096 * 4); cudaMalloc(&c, 8);
    const int iters = 5000;
    const char *names[] = {"FFMA2 column pairs, x shared by 4 rows", "scalar FFMA, same arithmetic", "FFMA2 row pairs, duplicated x", "FFMA2 column pairs, 8 accumulators"};
    for (int mode = 0; mode < 4; mode++)
        for (int nt : {128, 256}) {
            for (int rep = 0; rep < 2; rep++) {
                if (mode == 0) k<0><<<1, nt>>>(d, in, iters, c);
                if (mode == 1) k<1><<<1, nt>>>(d, in, iters, c);
                if (mode == 2) k<2><<<1, nt>>>(d, in, iters, c);
                if (mode == 3) k<3><<<1, nt>>>(d, in, iters, c);
                cudaDeviceSynchronize();
            }
            long long cy; cudaMemcpy(&cy, c, 8, cudaMemcpyDeviceToHost);
            printf("%-42s %d warps/SMSP: %7.1f clk per mat-vec per warp, %.2f clk per FFMA2-equivalent per SMSP\n", names[mode], nt / 128,
                   (double)cy / iters, (double)cy / ((double)iters * 84 * (nt / 128)));
        }
    return 0;
}
