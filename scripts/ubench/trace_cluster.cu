// Phase timeline of one clustered fused-kernel solve: clock64() marks of thread 0 of every CTA at the phase boundaries.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DGBDPCG_TRACE -I gbd_pcg_b200/csrc -I include \
//        [-DTR_NX=12 -DTR_N=512 -DTR_CL=16 -DTR_F64=0 -DTR_PMODE=P_REG] scripts/ubench/trace_cluster.cu -o scripts/ubench/trace_cluster
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "pcg_fused.cuh"
using namespace gbdpcg;
#ifndef TR_NX
#define TR_NX 12
#endif
#ifndef TR_N
#define TR_N 512
#endif
#ifndef TR_CL
#define TR_CL 16
#endif
#ifndef TR_F64
#define TR_F64 0
#endif
#ifndef TR_PMODE
#define TR_PMODE P_REG
#endif
#if TR_F64
typedef double TT;
#define TR_R 1
#else
typedef float TT;
#define TR_R 2
#endif

int main(int argc, char **argv)
{
    const int NX = TR_NX, N = TR_N, CL = TR_CL, iters = argc > 1 ? atoi(argv[1]) : 24;
    const int bpc = (N + CL - 1) / CL, threads = (bpc * (NX / TR_R) + 31) / 32 * 32;
    const size_t ms = 3ull * N * NX * NX, vs = (size_t)N * NX;
    std::vector<TT> S(ms, 0), P(ms, 0), g(vs);
    srand(1);
    auto rnd = [] { return (TT)rand() / RAND_MAX - (TT)0.5; };
    for (int k = 0; k < N; k++)
        for (int j = 0; j < 3; j++)
            for (int c = 0; c < NX; c++)
                for (int r = 0; r < NX; r++) {
                    TT v = (TT)0.02 * rnd();
                    if (j == 1 && r == c) v = -2;
                    S[((k * 3 + j) * NX + c) * NX + r] = v;
                    P[((k * 3 + j) * NX + c) * NX + r] = (j == 1 && r == c) ? (TT)-0.5 : (TT)0;
                }
    for (auto &x : g) x = rnd();
    TT *dS, *dP, *dg, *dl; uint32_t *dit;
    cudaMalloc(&dS, ms * sizeof(TT)); cudaMalloc(&dP, ms * sizeof(TT)); cudaMalloc(&dg, vs * sizeof(TT)); cudaMalloc(&dl, vs * sizeof(TT)); cudaMalloc(&dit, 4);
    cudaMemcpy(dS, S.data(), ms * sizeof(TT), cudaMemcpyHostToDevice); cudaMemcpy(dP, P.data(), ms * sizeof(TT), cudaMemcpyHostToDevice);
    cudaMemcpy(dg, g.data(), vs * sizeof(TT), cudaMemcpyHostToDevice);
    auto kern = pcg_fused_kernel<TT, NX, TR_R, TR_PMODE, true, true, 256>;
    const FusedPlan plan = make_fused_plan<TT>(NX, Geo<TT>::nxp(NX), bpc, TR_PMODE, TR_PMODE == P_REG ? 2 : 1);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.total);
    cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    KernelArgs<TT> a{};
    a.S = dS; a.Pinv = dP; a.gamma = dg; a.lambda = dl; a.iters = dit; a.exits = nullptr; a.mat_stride = ms; a.vec_stride = vs;
    a.N = N; a.batch = 1; a.bpc = bpc; a.max_iter = iters; a.two_buffers = (TR_PMODE == P_SMEM) ? 0 : 1; a.tol = 0;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(CL); lc.blockDim = dim3(threads); lc.dynamicSmemBytes = plan.total;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at; lc.numAttrs = 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f, msec = 0;
    for (int rep = 0; rep < 20; rep++) {
        cudaMemset(dl, 0, vs * sizeof(TT));
        cudaEventRecord(e0);
        void *args[] = {&a};
        cudaLaunchKernelExC(&lc, (const void *)kern, args);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&msec, e0, e1);
        if (msec < best) best = msec;
    }
    cudaError_t err = cudaDeviceSynchronize();
    printf("nx=%d N=%d %s cluster %d x %d threads, smem %u: status %s, %.2f us for %d iterations (%.0f clk/iter at 1.965 GHz)\n", NX, N,
           TR_F64 ? "f64" : "f32", CL, threads, plan.total, cudaGetErrorString(err), best * 1e3, iters, best * 1e-3 * 1.965e9 / iters);
#ifdef GBDPCG_TRACE
    static long long tr[16][64][8];
    cudaMemcpyFromSymbol(tr, g_trace, sizeof tr);
    const char *names[] = {"top", "pub r+BAR", "mv_P", "pub rt+BAR", "mv_S", "dots+all_sum", "scalars+upd+ghost"};
    for (int c : {0, CL / 2, CL - 1}) {
        printf("CTA %d: cycles per phase\n", c);
        for (int it = 8; it < 12 && it < iters; it++) {
            printf("  it %2d:", it);
            for (int m = 1; m <= 6; m++) printf(" %s=%lld", names[m], tr[c][it][m] - tr[c][it][m - 1]);
            printf(" | total %lld (to next top %lld)\n", tr[c][it][6] - tr[c][it][0], tr[c][it + 1][0] - tr[c][it][0]);
        }
    }
#endif
    return 0;
}
