// Phase timeline of the Schur-formation kernel (C2 shape, fp32): clock64 per phase, accumulated by thread 0 of every CTA.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DSCHUR_TRACE -I gbd_pcg_b200/csrc -I include \
//        scripts/ubench/trace_schur.cu -o scripts/ubench/trace_schur
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "kernel_table.h"
#include "schur_form.cuh"
using namespace gbdpcg;
int main(int argc, char **argv)
{
    const int NX = 14, nu = 7, N = 32, batch = argc > 1 ? atoi(argv[1]) : 8192, chunk = argc > 2 ? atoi(argv[2]) : 16;
    const size_t bb = NX * NX;
    std::vector<float> Q(batch * N * bb, 0.f), R(batch * (N - 1) * nu * nu, 0.f), A(batch * (N - 1) * bb), B(batch * (N - 1) * NX * nu), q(batch * N * NX, 1.f), r(batch * (N - 1) * nu, 1.f), c(batch * N * NX, 1.f);
    srand(1);
    for (auto &x : A) x = 0.1f * ((float)rand() / RAND_MAX - 0.5f);
    for (auto &x : B) x = 0.1f * ((float)rand() / RAND_MAX - 0.5f);
    for (size_t k = 0; k < (size_t)batch * N; k++) for (int i = 0; i < NX; i++) Q[k * bb + i * NX + i] = 2.f;
    for (size_t k = 0; k < (size_t)batch * (N - 1); k++) for (int i = 0; i < nu; i++) R[k * nu * nu + i * nu + i] = 0.5f;
    auto up = [](const std::vector<float> &h) { float *d; cudaMalloc(&d, h.size() * 4); cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice); return d; };
    SchurArgs a;
    a.Q = up(Q); a.R = up(R); a.A = up(A); a.B = up(B); a.q = up(q); a.r = up(r); a.c = up(c);
    float *S, *g, *Qi, *Ri; cudaMalloc(&S, (size_t)batch * 3 * N * bb * 4); cudaMalloc(&g, (size_t)batch * N * NX * 4); cudaMalloc(&Qi, (size_t)batch * N * bb * 4); cudaMalloc(&Ri, (size_t)batch * (N - 1) * nu * nu * 4);
    a.S = S; a.gamma = g; a.Qinv = Qi; a.Rinv = Ri; a.nx = NX; a.nu = nu; a.N = N; a.batch = batch;
    cudaMalloc(&a.trace, 64); cudaMemset(a.trace, 0, 64);
    auto kern = schur_form_nx_kernel<float, NX>;
    const size_t smem = schur_nx_smem_bytes<float>(NX, chunk, 256);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const uint32_t nchunks = (N + chunk - 1) / chunk;
    const int grid = 148 * (smem <= 112 * 1024 ? 2 : 1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaMemset(a.trace, 0, 64);
        cudaEventRecord(e0);
        kern<<<grid, 256, smem>>>(a, (uint32_t)chunk, nchunks);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    }
    unsigned long long tr[8]; cudaMemcpy(tr, a.trace, 64, cudaMemcpyDeviceToHost);
    printf("schur_form_nx<f32,14> batch %d chunk %d smem %zu grid %d: %.1f us (%s)\n", batch, chunk, smem, grid, ms * 1e3, cudaGetErrorString(cudaGetLastError()));
    const char *names[] = {"stage", "invert", "AQ/BR", "diag+gamma", "store"};
    const double items = (double)batch * nchunks;
    for (int i = 0; i < 5; i++) printf("  %-11s %8.0f clk per chunk\n", names[i], tr[i] / items);
    return 0;
}
