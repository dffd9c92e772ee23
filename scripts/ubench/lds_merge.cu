// Micro-benchmark: how many shared-memory wavefronts does an LDS.{32,64,128} cost when lanes of a warp read a
// few DISTINCT addresses (broadcast groups)?  Decides the window layout of the PCG kernel.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a lds_merge.cu -o lds_merge && ./lds_merge
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int W>  // bytes per lane: 4, 8, 16
__global__ void k(const int *lane_off, long long *out, int iters, float *sink)
{
    extern __shared__ __align__(16) unsigned char sm[];
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = (float)i;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + lane_off[lane];
    float accv[16];
#pragma unroll
    for (int u = 0; u < 16; u++) accv[u] = 0.f;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            unsigned a = base + u * 1024 + ((it & 1) << 14);  // 16 independent loads; address varies with it
            if constexpr (W == 16) {
                float x, y, z, w;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a));
                accv[u] = __int_as_float(__float_as_int(accv[u]) ^ __float_as_int(x));
            } else if constexpr (W == 8) {
                float x, y;
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(a));
                accv[u] = __int_as_float(__float_as_int(accv[u]) ^ __float_as_int(x));
            } else {
                float x;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a));
                accv[u] = __int_as_float(__float_as_int(accv[u]) ^ __float_as_int(x));
            }
        }
    }
    long long t1 = clock64();
    if (lane == 0) out[threadIdx.x >> 5] = t1 - t0;
    float acc = 0.f;
#pragma unroll
    for (int u = 0; u < 16; u++) acc += accv[u];
    if (acc == 123.456f) *sink = acc;
}

struct Pat { const char *name; int off[32]; };

int main()
{
    cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); cudaFuncSetAttribute(k<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); cudaFuncSetAttribute(k<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    int *d_off; long long *d_out; float *d_sink;
    cudaMalloc(&d_off, 32 * 4); cudaMalloc(&d_out, 64 * 8); cudaMalloc(&d_sink, 4);
    auto groups14 = [](int spacing, Pat &p) { for (int l = 0; l < 32; l++) p.off[l] = (l / 14) * spacing; };
    Pat pats[20]; int np = 0;
    { Pat p{"all lanes same address"}; for (int l = 0; l < 32; l++) p.off[l] = 0; pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 80 B apart (current NXP=20)"}; groups14(80, p); pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 64 B apart"}; groups14(64, p); pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 16 B apart (same 128 B line)"}; groups14(16, p); pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 32 B apart"}; groups14(32, p); pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 128 B apart (same banks)"}; groups14(128, p); pats[np++] = p; }
    { Pat p{"3 groups of 14 lanes, 144 B apart"}; groups14(144, p); pats[np++] = p; }
    { Pat p{"2 groups of 16 lanes, 16 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 16) * 16; pats[np++] = p; }
    { Pat p{"2 groups of 16 lanes, 80 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 16) * 80; pats[np++] = p; }
    { Pat p{"4 groups of 8 lanes, 16 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 8) * 16; pats[np++] = p; }
    { Pat p{"4 groups of 8 lanes, 80 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 8) * 80; pats[np++] = p; }
    { Pat p{"groups of 7 lanes (5 groups), 16 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 7) * 16; pats[np++] = p; }
    { Pat p{"groups of 7 lanes (5 groups), 80 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 7) * 80; pats[np++] = p; }
    { Pat p{"8 groups of 4 lanes, 16 B apart (one full line)"}; for (int l = 0; l < 32; l++) p.off[l] = (l / 4) * 16; pats[np++] = p; }
    { Pat p{"every lane distinct, 16 B apart (512 B)"}; for (int l = 0; l < 32; l++) p.off[l] = l * 16; pats[np++] = p; }
    { Pat p{"every lane distinct, 4 B apart (LDS.32 ideal)"}; for (int l = 0; l < 32; l++) p.off[l] = l * 4; pats[np++] = p; }
    { Pat p{"every lane distinct, 8 B apart (LDS.64 ideal)"}; for (int l = 0; l < 32; l++) p.off[l] = l * 8; pats[np++] = p; }
    { Pat p{"interleaved groups: lane%3 -> group, 16 B apart"}; for (int l = 0; l < 32; l++) p.off[l] = (l % 3) * 16; pats[np++] = p; }

    const int iters = 100000;
    printf("%-58s %8s %8s %8s   (ns per LDS instruction SM-wide, CUDA events; at 1.965 GHz 1 cycle = 0.509 ns)\n", "pattern", "W=4", "W=8", "W=16");
    for (int nw : {1, 4, 16}) {
        printf("-- %d warp(s)\n", nw);
        for (int i = 0; i < np; i++) {
            double r[3];
            for (int wi = 0; wi < 3; wi++) {
                // make offsets legal for the width (alignment) -- all offsets are multiples of 4; widen to alignment
                int off[32];
                const int W = wi == 0 ? 4 : wi == 1 ? 8 : 16;
                bool ok = true;
                for (int l = 0; l < 32; l++) { off[l] = pats[i].off[l]; if (off[l] % W) ok = false; }
                if (!ok) { r[wi] = -1; continue; }
                cudaMemcpy(d_off, off, sizeof off, cudaMemcpyHostToDevice);
                cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
                float ms = 0;
                for (int rep = 0; rep < 2; rep++) {
                    cudaEventRecord(e0);
                    if (wi == 0) k<4><<<1, 32 * nw, 65536>>>(d_off, d_out, iters, d_sink);
                    if (wi == 1) k<8><<<1, 32 * nw, 65536>>>(d_off, d_out, iters, d_sink);
                    if (wi == 2) k<16><<<1, 32 * nw, 65536>>>(d_off, d_out, iters, d_sink);
                    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
                }
                // ns per LDS instruction SM-wide: total instructions = nw * iters * 16
                r[wi] = (double)ms * 1e6 / (iters * 16.0) / nw;
            }
            printf("%-58s %8.2f %8.2f %8.2f\n", pats[i].name, r[0], r[1], r[2]);
        }
    }
    printf("(16-warp rows: cycles per LDS instruction SM-wide = wavefronts per instruction when the LSU data pipe is the limit)\n");
    return 0;
}
