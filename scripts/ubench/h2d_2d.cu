// Micro-benchmark: pinned host -> device bandwidth of the symmetric-S shipping pattern (cudaMemcpy2DAsync: 2*nx^2 of every 3*nx^2
// elements, nx = 14 -> rows of 1568 B at a pitch of 2352 B) against one contiguous copy of the same bytes, on 1, 2 and 4 streams.
//   nvcc -O3 -std=c++17 scripts/ubench/h2d_2d.cu -o scripts/ubench/h2d_2d
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
int main()
{
    const size_t rows = 8192ull * 32, pitch = 2352, width = 1568;
    char *h, *d;
    cudaHostAlloc(&h, rows * pitch, cudaHostAllocDefault);
    cudaMalloc(&d, rows * pitch);
    for (size_t i = 0; i < rows * pitch; i += 4096) h[i] = (char)i;
    cudaStream_t st[8];
    for (auto &s : st) cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto run = [&](const char *name, int nstreams, int pieces, bool twoD, size_t bytes_moved) {
        float best = 1e9f;
        for (int rep = 0; rep < 5; rep++) {
            cudaDeviceSynchronize();
            cudaEventRecord(e0, 0);
            for (int s = 0; s < nstreams; s++) cudaStreamWaitEvent(st[s], e0, 0);
            for (int p = 0; p < pieces; p++) {
                const size_t r0 = rows * p / pieces, r1 = rows * (p + 1) / pieces;
                if (twoD) cudaMemcpy2DAsync(d + r0 * pitch + 784, pitch, h + r0 * pitch + 784, pitch, width, r1 - r0, cudaMemcpyHostToDevice, st[p % nstreams]);
                else cudaMemcpyAsync(d + r0 * width, h + r0 * width, (r1 - r0) * width, cudaMemcpyHostToDevice, st[p % nstreams]);
            }
            std::vector<cudaEvent_t> done(nstreams);
            for (int s = 0; s < nstreams; s++) { cudaEventCreate(&done[s]); cudaEventRecord(done[s], st[s]); cudaStreamWaitEvent(0, done[s], 0); }
            cudaEventRecord(e1, 0);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
            for (auto &e : done) cudaEventDestroy(e);
        }
        printf("%-44s streams %d pieces %2d: %7.3f ms  %6.2f GB/s  (%s)\n", name, nstreams, pieces, best, bytes_moved / best / 1e6, cudaGetErrorString(cudaGetLastError()));
    };
    const size_t b2 = rows * width;
    run("contiguous, 2/3 of S worth of bytes", 1, 1, false, b2);
    run("contiguous", 1, 8, false, b2);
    run("contiguous", 2, 8, false, b2);
    run("2-D 1568 of 2352 B", 1, 1, true, b2);
    run("2-D 1568 of 2352 B", 1, 8, true, b2);
    run("2-D 1568 of 2352 B", 2, 8, true, b2);
    run("2-D 1568 of 2352 B", 4, 16, true, b2);
    run("2-D 1568 of 2352 B", 8, 32, true, b2);
    run("contiguous full S (3/3)", 1, 1, false, rows * pitch * 0 + b2);   // same bytes, reference line
    return 0;
}
