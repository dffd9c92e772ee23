// Phase timeline of one fused-kernel solve (C2 shape: nx=14, N=32, fp32, Phi^-1 in registers): clock64() marks
// from lane 0 of every warp at the phase boundaries of each iteration.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DGBDPCG_TRACE -I gbd_pcg_b200/csrc -I include \
//        scripts/ubench/trace_fused.cu -o scripts/ubench/trace_fused
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "pcg_fused.cuh"
using namespace gbdpcg;

int main(int argc, char **argv)
{
    #ifndef TR_N
#define TR_N 32
#endif
    const int NX = 14, N = TR_N, iters = argc > 1 ? atoi(argv[1]) : 24, batch = argc > 2 ? atoi(argv[2]) : 1;
    const size_t ms = 3ull * N * NX * NX, vs = (size_t)N * NX;
    std::vector<float> S(ms * batch, 0.f), P(ms * batch, 0.f), g(vs * batch);
    srand(1);
    auto rnd = [] { return (float)rand() / RAND_MAX - 0.5f; };
    for (int pb = 0; pb < batch; pb++)
        for (int k = 0; k < N; k++)
            for (int j = 0; j < 3; j++)
                for (int c = 0; c < NX; c++)
                    for (int r = 0; r < NX; r++) {
                        float v = 0.02f * rnd();
                        if (j == 1 && r == c) v = -2.f;
                        S[pb * ms + ((k * 3 + j) * NX + c) * NX + r] = v;
                        P[pb * ms + ((k * 3 + j) * NX + c) * NX + r] = (j == 1 && r == c) ? -0.5f : 0.f;
                    }
    for (auto &x : g) x = rnd();
    float *dS, *dP, *dg, *dl; uint32_t *dit;
    cudaMalloc(&dS, S.size() * 4); cudaMalloc(&dP, P.size() * 4); cudaMalloc(&dg, g.size() * 4); cudaMalloc(&dl, g.size() * 4); cudaMalloc(&dit, 4 * batch);
    cudaMemcpy(dS, S.data(), S.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dP, P.data(), P.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dg, g.data(), g.size() * 4, cudaMemcpyHostToDevice);
#ifndef TR_R
#define TR_R 2
#endif
#ifndef TR_MAXT
#define TR_MAXT 256
#endif
    auto kern = pcg_fused_kernel<float, NX, TR_R, P_REG, true, false, TR_MAXT>;
    const FusedPlan plan = make_fused_plan<float>(NX, Geo<float>::nxp(NX), N, P_REG, 2);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total);
    KernelArgs<float> a{};
    a.S = dS; a.Pinv = dP; a.gamma = dg; a.lambda = dl; a.iters = dit; a.exits = nullptr; a.mat_stride = ms; a.vec_stride = vs;
    a.N = N; a.batch = batch; a.bpc = N; a.max_iter = iters; a.two_buffers = 1; a.tol = 0.f;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float msec = 0;
    float best = 1e9f;
    for (int rep = 0; rep < 20; rep++) {
        cudaMemset(dl, 0, g.size() * 4);
        cudaEventRecord(e0);
        kern<<<batch < 148 ? batch : 148, N * NX / TR_R, plan.total>>>(a);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&msec, e0, e1);
        if (msec < best) best = msec;
    }
    msec = best;
    cudaError_t err = cudaDeviceSynchronize();
    printf("status %s, kernel %.2f us for %d iterations x %d systems\n", cudaGetErrorString(err), msec * 1e3, iters, batch);
#ifndef GBDPCG_TRACE
    return 0;
#else
    static long long tr[16][64][8], fx[16][8];
    cudaMemcpyFromSymbol(tr, g_trace, sizeof tr);
    cudaMemcpyFromSymbol(fx, g_trace_fix, sizeof fx);
    const char *names[] = {"top", "pub r+BAR", "mv_P", "pub rt+BAR", "mv_S", "dots+all_sum", "scalars+upd"};
    for (int it = 8; it < 12 && it < iters; it++) {
        printf("  it %2d:", it);
        for (int m = 1; m <= 6; m++) printf(" %s=%lld", names[m], tr[0][it][m] - tr[0][it][m - 1]);
        printf(" | total %lld (to next top %lld)\n", tr[0][it][6] - tr[0][it][0], tr[0][it + 1][0] - tr[0][it][0]);
    }
    // fixed costs of the LAST system this CTA solved (thread 0): clocks from kernel entry
    printf("fixed: init done +%lld, matrices arrived +%lld, rows in registers +%lld, first iteration top +%lld, loop left +%lld, kernel end +%lld\n",
           fx[0][1] - fx[0][0], fx[0][2] - fx[0][0], fx[0][3] - fx[0][0], tr[0][0][0] - fx[0][0], fx[0][5] - fx[0][0], fx[0][6] - fx[0][0]);
    return 0;
#endif
}
