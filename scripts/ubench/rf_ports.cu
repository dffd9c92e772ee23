// Micro-benchmark: what bounds a register-resident mat-vec on sm_100a -- the FP32 pipe or the register file's ports?
// Every thread keeps 84 packed matrix pairs (one mat-vec's worth of pcg_tmem4.cuh: 4 rows x 42 columns) in registers and streams
// them through fma.rn.f32x2 / fma.rn.f32 under different operand arrangements.  The multiply-adds are asm volatile, so the SASS
// stream is exactly what is written here (no CSE, no dead-code removal); ptxas still picks the registers.
//   A  FFMA2  acc[j] = m[i] * x + acc[j]        m unique, x shared by 4 consecutive instructions, 4 rotating accumulators (the kernel)
//   B  FFMA2  acc[j] = m[i] * m[i+1] + acc[j]   every source operand unique
//   C  FFMA2  acc[j] = m[i] * x[k] + acc[j]     x rotates over three pairs from instruction to instruction (no reuse)
//   D  FFMA2  acc[j] = x * x + acc[j]           no unique source: accumulator read + write only
//   E  FFMA   acc[j] = m[i] * x + acc[j]        scalar form of A (168 instructions)
//   F  FFMA   acc[j] = m[i] * m[i+1] + acc[j]   scalar form of B
//   G  FFMA   acc[j] = x * x + acc[j]           scalar form of D
// (Forms that only WRITE a result -- d[j] = m[i] * x + z -- cannot be timed: ptxas removes the overwritten results whatever the
// inline assembly says.)
// Prints clocks per instruction per SM sub-partition for 1 and 2 warps per sub-partition, and the 32-bit register-file words an
// instruction moves (reads of non-reused operands + the write) -- the model "two words per clock per lane, reads and writes together"
// predicts clocks = words / 2 (never below the pipe's 2 clk per FFMA2 / 1 clk per FFMA).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/ubench/rf_ports.cu -o scripts/ubench/rf_ports
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define FMA2(d, a, b, c) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c) : "memory")
#define FMA1(d, a, b, c) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c) : "memory")

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float *out, const unsigned long long *in, const float *inf, int iters, long long *cyc)
{
    unsigned long long m[84], acc[4], x[3], z;
    float ms[168], accs[8], xs[6], zs;
#pragma unroll
    for (int i = 0; i < 84; i++) m[i] = in[threadIdx.x + 512 * i];
#pragma unroll
    for (int i = 0; i < 168; i++) ms[i] = (MODE == 4 || MODE == 5) ? inf[threadIdx.x + 512 * i] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) acc[i] = 0ull;
#pragma unroll
    for (int i = 0; i < 8; i++) accs[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) x[i] = in[threadIdx.x + 7 + i];
#pragma unroll
    for (int i = 0; i < 6; i++) xs[i] = inf[threadIdx.x + 11 + i];
    z = in[threadIdx.x + 3];
    zs = inf[threadIdx.x + 5];
    // opaque to the compiler from here on: the values must stay in registers (no re-loading inside the timed loop)
#pragma unroll
    for (int i = 0; i < 84; i++) if (MODE <= 2) asm volatile("" : "+l"(m[i]));
    if constexpr (MODE == 4 || MODE == 5) {
#pragma unroll
        for (int i = 0; i < 168; i++) asm volatile("" : "+f"(ms[i]));
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if constexpr (MODE <= 3) {
#pragma unroll
            for (int q = 0; q < 21; q++) {
#pragma unroll
                for (int row = 0; row < 4; row++) {
                    const int i = 4 * q + row;
                    if constexpr (MODE == 0) FMA2(acc[row], m[i], x[q % 3], acc[row]);
                    if constexpr (MODE == 1) FMA2(acc[row], m[i], m[(i + 1) % 84], acc[row]);
                    if constexpr (MODE == 2) FMA2(acc[row], m[i], x[i % 3], acc[row]);
                    if constexpr (MODE == 3) FMA2(acc[row], x[q % 3], x[q % 3], acc[row]);
                }
            }
        } else {
#pragma unroll
            for (int q = 0; q < 42; q++) {
#pragma unroll
                for (int row = 0; row < 4; row++) {
                    const int i = 4 * q + row;
                    if constexpr (MODE == 4) FMA1(accs[row + 4 * (q & 1)], ms[i], xs[q % 6], accs[row + 4 * (q & 1)]);
                    if constexpr (MODE == 5) FMA1(accs[row + 4 * (q & 1)], ms[i], ms[(i + 1) % 168], accs[row + 4 * (q & 1)]);
                    if constexpr (MODE == 6) FMA1(accs[row + 4 * (q & 1)], xs[q % 6], xs[q % 6], accs[row + 4 * (q & 1)]);
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) s += __uint_as_float((unsigned)acc[i]) + __uint_as_float((unsigned)(acc[i] >> 32));
#pragma unroll
    for (int i = 0; i < 8; i++) s += accs[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main()
{
    float *d, *inf;
    unsigned long long *in;
    long long *c;
    cudaMalloc(&d, 4096 * 4);
    cudaMalloc(&in, (512 * 90) * 8);
    cudaMalloc(&inf, (512 * 180) * 4);
    cudaMemset(in, 0, (512 * 90) * 8);
    cudaMemset(inf, 0, (512 * 180) * 4);
    cudaMalloc(&c, 8);
    const int iters = 4000;
    const char *names[] = {"A FFMA2 acc[j] = m[i]*x + acc[j]      (the kernel's pattern)", "B FFMA2 acc[j] = m[i]*m[i+1] + acc[j] (all sources unique)",
                           "C FFMA2 acc[j] = m[i]*x[k] + acc[j]   (x not reusable)", "D FFMA2 acc[j] = x*x + acc[j]         (accumulator read + write only)",
                           "E FFMA  acc[j] = m[i]*x + acc[j]      (scalar form of A)", "F FFMA  acc[j] = m[i]*m[i+1] + acc[j] (scalar form of B)",
                           "G FFMA  acc[j] = x*x + acc[j]         (scalar form of D)"};
    const float words[] = {2 + 0.5f + 2 + 2, 2 + 2 + 2 + 2, 2 + 2 + 2 + 2, 2 + 2, 1 + 0.25f + 1 + 1, 1 + 1 + 1 + 1, 1 + 1};
    const float pipe[] = {2, 2, 2, 2, 1, 1, 1};
    for (int mode = 0; mode < 7; mode++)
        for (int nt : {128, 256}) {
            long long best = 1ll << 62;
            for (int rep = 0; rep < 3; rep++) {
                if (mode == 0) k<0><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 1) k<1><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 2) k<2><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 3) k<3><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 4) k<4><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 5) k<5><<<1, nt>>>(d, in, inf, iters, c);
                if (mode == 6) k<6><<<1, nt>>>(d, in, inf, iters, c);
                long long h = 0;
                cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
                if (h < best) best = h;
            }
            const int wps = nt / 128, ninst = (mode >= 4) ? 168 : 84;
            const double clk = (double)best / iters / ninst / wps;
            const double model = words[mode] / 2 > pipe[mode] ? words[mode] / 2 : pipe[mode];
            printf("%-72s %d warps/SMSP: %5.2f clk per instruction per SMSP | %.2f register-file words -> model %.2f clk | %s\n", names[mode], wps, clk,
                   words[mode], model, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
