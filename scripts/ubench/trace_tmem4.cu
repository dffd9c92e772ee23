// Phase timeline of the four-rows-per-thread tensor-memory kernel (pcg_tmem4.cuh) on a full batch: C2 shape (nx 14, N 32, fp32),
// 8192 systems on 296 CTAs (two per SM), a fixed number of iterations per system (exit tolerance 0).  clock64() marks of thread 0
// of the first CTAs: per system where the time between two systems goes (waits for the bulk copies, staging -> tensor memory /
// registers), and inside one system the phases of an iteration.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DTMEM4_TRACE -I gbd_pcg_b200/csrc -I include \
//        scripts/ubench/trace_tmem4.cu -o scripts/ubench/trace_tmem4
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "pcg_tmem4.cuh"
using namespace gbdpcg;

#ifndef TR_SQT
#define TR_SQT 2
#endif
#ifndef TR_PQR
#define TR_PQR 0
#endif

int main(int argc, char **argv)
{
    const int NX = 14, N = 32, iters = argc > 1 ? atoi(argv[1]) : 25, batch = argc > 2 ? atoi(argv[2]) : 8192;
    const int grid = argc > 3 ? atoi(argv[3]) : 296;
    const int solo = argc > 4 ? atoi(argv[4]) : 0;          // 1: the solo form (second staging image, at most one CTA per SM)
    const size_t ms = 3ull * N * NX * NX, vs = (size_t)N * NX;
    std::vector<float> S(ms, 0.f), P(ms, 0.f), g(vs);
    srand(1);
    auto rnd = [] { return (float)rand() / RAND_MAX - 0.5f; };
    for (int k = 0; k < N; k++)
        for (int j = 0; j < 3; j++)
            for (int c = 0; c < NX; c++)
                for (int r = 0; r < NX; r++) {
                    float v = 0.02f * rnd();
                    if (j == 1 && r == c) v = -2.f;
                    S[((k * 3 + j) * NX + c) * NX + r] = v;
                    P[((k * 3 + j) * NX + c) * NX + r] = (j == 1 && r == c) ? -0.5f : 0.f;
                }
    for (auto &x : g) x = rnd();
    float *dS, *dP, *dg, *dl; uint32_t *dit;
    cudaMalloc(&dS, ms * 4 * batch); cudaMalloc(&dP, ms * 4 * batch); cudaMalloc(&dg, vs * 4 * batch); cudaMalloc(&dl, vs * 4 * batch); cudaMalloc(&dit, 4 * batch);
    for (int b = 0; b < batch; b++) {        // every system its own copy in HBM (same values): the traffic is real
        cudaMemcpy(dS + b * ms, S.data(), ms * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dP + b * ms, P.data(), ms * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dg + b * vs, g.data(), vs * 4, cudaMemcpyHostToDevice);
    }
    auto kern = pcg_tmem4_kernel<NX, TR_SQT, TR_PQR>;
    const Tmem4Plan plan = make_tmem4_plan(NX, t4_nxp(NX), N);
    const size_t smem = plan.total + (solo ? t4_stage_bytes(NX, N) : 0);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    KernelArgs<float> a{};
    a.S = dS; a.Pinv = dP; a.gamma = dg; a.lambda = dl; a.iters = dit; a.exits = nullptr; a.mat_stride = ms; a.vec_stride = vs;
    a.N = N; a.batch = batch; a.bpc = N; a.max_iter = iters; a.two_buffers = solo; a.use_tma = 1; a.tol = 0.f;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float msec = 0, best = 1e9f;
    for (int rep = 0; rep < 10; rep++) {
        cudaMemset(dl, 0, vs * 4 * batch);
        cudaEventRecord(e0);
        kern<<<grid, 128, smem>>>(a);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&msec, e0, e1);
        if (msec < best) best = msec;
    }
    cudaError_t err = cudaDeviceSynchronize();
    printf("status %s, kernel %.1f us for %d systems x %d iterations on %d CTAs (S_tmem=%d, P_reg=%d)\n", cudaGetErrorString(err), best * 1e3, batch, iters, grid,
           TR_SQT, TR_PQR);
#ifdef TMEM4_TRACE
    static long long sy[8][40][8], itr[8][48][5];
    cudaMemcpyFromSymbol(sy, g_t4sys, sizeof sy);
    cudaMemcpyFromSymbol(itr, g_t4it, sizeof itr);
    const int nsys = (batch + grid - 1) / grid < 40 ? (batch + grid - 1) / grid : 40;
    {
        const char *fn[] = {"wait Pinv", "Pinv->TMEM", "wait S", "S->regs", "prologue", "iterations", "outputs"};
        printf("CTA 0, FIRST system, clocks from kernel entry: loop top +%lld;", sy[0][0][0] - itr[0][47][0]);
        for (int m = 0; m < 7; m++) printf(" %s=%lld", fn[m], sy[0][0][m + 1] - sy[0][0][m]);
        printf(" | kernel end +%lld after the outputs\n", itr[0][47][1] - sy[0][0][7]);
    }
    const char *sn[] = {"wait Pinv", "Pinv->TMEM", "wait S", "S->regs", "prologue", "iterations", "outputs"};
    for (int c = 0; c < 2; c++) {
        double sum[7] = {0, 0, 0, 0, 0, 0, 0};
        int cnt = 0;
        for (int k = 2; k + 1 < nsys; k++, cnt++)
            for (int m = 0; m < 7; m++) sum[m] += (double)(sy[c][k][m + 1] - sy[c][k][m]);
        printf("CTA %d, mean over systems 2..%d (clocks):", c, nsys - 2);
        double tot = 0;
        for (int m = 0; m < 7; m++) { printf(" %s=%.0f", sn[m], sum[m] / cnt); tot += sum[m] / cnt; }
        printf(" | system total %.0f, of which outside the iterations %.0f (%.1f %%)\n", tot, tot - sum[5] / cnt, 100 * (tot - sum[5] / cnt) / tot);
    }
    const char *in[] = {"pub r+BAR+r~", "pub r~+BAR+w", "dots+reduction", "scalars+updates"};
    for (int c = 0; c < 2; c++)
        for (int it = 8; it < 12 && it < iters; it++) {
            printf("  CTA %d it %2d:", c, it);
            for (int m = 0; m < 4; m++) printf(" %s=%lld", in[m], itr[c][it][m + 1] - itr[c][it][m]);
            printf(" | iteration %lld\n", itr[c][it + 1][0] - itr[c][it][0]);
        }
#endif
    return 0;
}
