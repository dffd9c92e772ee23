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
    static long long tr[8][64][16];
    cudaMemcpyFromSymbol(tr, g_trace, sizeof tr);
    const char *names[] = {"top", "pub r + BAR", "mv_P", "pub rt + BAR", "mv_S", "dots+shfl", "STS+BAR", "tree", "scalars+upd"};
    for (int w : {0, TR_N * 7 / 32 - 1}) {
        printf("warp %d: cycles per phase, iterations 8..11\n", w);
        for (int it = 8; it < 12 && it < iters; it++) {
            printf("  it %2d:", it);
            for (int m = 1; m <= 8; m++) printf(" %s=%lld", names[m], tr[w][it][m] - tr[w][it][m - 1]);
            printf(" | total %lld (to next top %lld)\n", tr[w][it][8] - tr[w][it][0], tr[w][it + 1][0] - tr[w][it][0]);
        }
    }
    long long span = tr[0][iters - 1][8] - tr[0][0][0];
    printf("warp 0: %lld cycles for %d iterations = %.1f cycles/iteration\n", span, iters, (double)span / iters);
    return 0;
#endif
}
