/*
 * TEST / BASELINE INFRASTRUCTURE ONLY.  Thin extern "C" doors onto the UNMODIFIED reference
 * (A2R-Lab/GBD-PCG @ 75d6214), whose headers are #included from where they lie
 * (-I/root/reference/include -I/root/reference/GLASS, see oracle/Makefile).  Nothing from the
 * reference is copied here.  One shared object is built per (T, STATE_SIZE, KNOT_POINTS)
 * because the reference instantiates its kernel from those macros (include/interface.cuh:110)
 * and defines mutable globals in a header (include/constants.cuh:15-19).
 *
 *   -DREF_T=float|double -DSTATE_SIZE=nx -DKNOT_POINTS=N
 *
 * ref_glass_build_pinv: glass::invertMatrix / glass::gemm (the reference's primitives for the preconditioner) on the blocks of S.
 * Three doors = the three timing scopes of SURVEY.md s8(d):
 *   ref_kernel_launch : direct cudaLaunchCooperativeKernel of pcg<T,nx,N>, as include/interface.cuh:130-132
 *   ref_solve_dev     : the reference's device-pointer solvePCG overload (include/interface.cuh:92-144)
 *   ref_solve_host    : the reference's host-pointer solvePCG overload (include/interface.cuh:24-89);
 *                       time only: it never initialises d_Pinv, so its result is undefined.
 */
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "gpu_pcg.cuh"

#ifndef REF_T
#define REF_T float
#endif
typedef REF_T T;

extern "C" {

void ref_dims(uint32_t *nx, uint32_t *N, int *is_f64)
{
    *nx = STATE_SIZE;
    *N = KNOT_POINTS;
    *is_f64 = (sizeof(T) == 8);
}

/* smem the reference would request, plus what its kernel really touches. */
void ref_smem_bytes(uint64_t *requested, uint64_t *used)
{
    const uint64_t nx = STATE_SIZE, N = KNOT_POINTS, M = (N > nx ? N : nx);
    *requested = pcgSharedMemSize<T>(STATE_SIZE, KNOT_POINTS);
    *used = sizeof(T) * (6 * nx * nx + 12 * nx + 2 * M);
}

/* Scope K.  Asynchronous on the legacy default stream; d_iters/d_exit are caller-owned device
 * scalars.  extra_smem lets the harness cover the reference's 2*nx-element under-allocation
 *  without editing the reference.  Returns the cudaError_t of the launch. */
int ref_kernel_launch(void *d_S, void *d_Pinv, void *d_gamma, void *d_lambda, void *d_r, void *d_p,
                      void *d_v_temp, void *d_eta_new_temp, uint32_t *d_iters, bool *d_exit,
                      uint32_t max_iter, double exit_tol, uint64_t extra_smem)
{
    void *kernel = (void *)pcg<T, STATE_SIZE, KNOT_POINTS>;
    T tol = (T)exit_tol;
    void *args[] = {&d_S, &d_Pinv, &d_gamma, &d_lambda, &d_r, &d_p, &d_v_temp, &d_eta_new_temp,
                    &d_iters, &d_exit, &max_iter, &tol};
    size_t smem = pcgSharedMemSize<T>(STATE_SIZE, KNOT_POINTS) + extra_smem;
    cudaError_t e = cudaLaunchCooperativeKernel(kernel, dim3(KNOT_POINTS), pcg_constants::DEFAULT_BLOCK, args, smem);
    if (e == cudaSuccess) e = cudaPeekAtLastError();
    return (int)e;
}

/* Same launch on a caller-chosen stream with caller-owned scratch: lets the bench keep several reference solves in
 * flight at once (the reference's own entry points use the legacy default stream only, include/interface.cuh:132). */
int ref_kernel_launch_stream(void *d_S, void *d_Pinv, void *d_gamma, void *d_lambda, void *d_r, void *d_p,
                             void *d_v_temp, void *d_eta_new_temp, uint32_t *d_iters, bool *d_exit,
                             uint32_t max_iter, double exit_tol, uint64_t extra_smem, void *stream)
{
    void *kernel = (void *)pcg<T, STATE_SIZE, KNOT_POINTS>;
    T tol = (T)exit_tol;
    void *args[] = {&d_S, &d_Pinv, &d_gamma, &d_lambda, &d_r, &d_p, &d_v_temp, &d_eta_new_temp,
                    &d_iters, &d_exit, &max_iter, &tol};
    size_t smem = pcgSharedMemSize<T>(STATE_SIZE, KNOT_POINTS) + extra_smem;
    cudaError_t e = cudaLaunchCooperativeKernel(kernel, dim3(KNOT_POINTS), pcg_constants::DEFAULT_BLOCK, args, smem,
                                                (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaPeekAtLastError();
    return (int)e;
}

/* Scope D.  Blocking (the reference reads the iteration count back with cudaMemcpy). */
uint32_t ref_solve_dev(void *d_S, void *d_Pinv, void *d_gamma, void *d_lambda, void *d_r, void *d_p,
                       void *d_v_temp, void *d_eta_new_temp, uint32_t max_iter, double exit_tol)
{
    pcg_config<T> cfg;
    cfg.pcg_exit_tol = (T)exit_tol;
    cfg.pcg_max_iter = max_iter;
    return solvePCG<T>((uint32_t)STATE_SIZE, (uint32_t)KNOT_POINTS, (T *)d_S, (T *)d_Pinv, (T *)d_gamma,
                       (T *)d_lambda, (T *)d_r, (T *)d_p, (T *)d_v_temp, (T *)d_eta_new_temp, &cfg);
}

/* Scope H.  Blocking; h_S and h_gamma are written back unchanged by the reference. */
uint32_t ref_solve_host(void *h_S, void *h_gamma, void *h_lambda, uint32_t max_iter, double exit_tol)
{
    pcg_config<T> cfg;
    cfg.pcg_exit_tol = (T)exit_tol;
    cfg.pcg_max_iter = max_iter;
    return solvePCG<T>((T *)h_S, (T *)h_gamma, (T *)h_lambda, (unsigned)STATE_SIZE, (unsigned)KNOT_POINTS, &cfg);
}

/* ---- the reference's own primitives for the preconditioner (GLASS/src/L3/inv.cuh:6-37 glass::invertMatrix, GLASS/src/L3/gemm.cuh
 * glass::gemm; carried by the reference tree, never called on its path): Phi^-1 of ONE system in the solver's layout, one CTA per
 * block-row.  kind 1: Phi^-1_kk = S_kk^-1 ; kind 2: + Phi^-1_{k,k+-1} = -S_kk^-1 (S_{k,k+-1} S_{k+-1,k+-1}^-1).  Only the data
 * movement around the two primitives is harness code; pins oracle_build_pinv_* and the product's pinv_build kernel. ---- */
__global__ void ref_glass_pinv_kernel(const T *S, T *P, int kind)
{
    constexpr unsigned nx = STATE_SIZE, N = KNOT_POINTS, bb = nx * nx;
    extern __shared__ unsigned char ref_smem_raw[];
    T *aug = (T *)ref_smem_raw;          /* 3 augmented images [A | I], nx x 2nx each: block-rows k-1, k, k+1 */
    T *s_temp = aug + 3 * 2 * bb;        /* 2nx + 1 */
    T *tmp = s_temp + 2 * nx + 1;        /* nx x nx */
    T *out = tmp + bb;                   /* nx x nx */
    const unsigned k = blockIdx.x;
    for (unsigned i = threadIdx.x; i < 3 * bb; i += blockDim.x) P[(size_t)k * 3 * bb + i] = (T)0;
    __syncthreads();
    for (int j = -1; j <= 1; j++) {
        if ((int)k + j < 0 || k + j >= N) continue;
        if (j != 0 && kind != 2) continue;
        T *A = aug + (j + 1) * 2 * bb;
        for (unsigned i = threadIdx.x; i < bb; i += blockDim.x) {
            A[i] = S[(size_t)(k + j) * 3 * bb + bb + i];
            A[bb + i] = (i / nx == i % nx) ? (T)1 : (T)0;
        }
        __syncthreads();
        glass::invertMatrix<T>(nx, A, s_temp);      /* leaves [I | A^-1] */
        __syncthreads();
    }
    const T *Dk = aug + 1 * 2 * bb + bb;
    for (unsigned i = threadIdx.x; i < bb; i += blockDim.x) P[(size_t)k * 3 * bb + bb + i] = Dk[i];
    if (kind == 2) {
        for (int side = 0; side < 2; side++) {
            if ((side == 0 && k == 0) || (side == 1 && k == N - 1)) continue;
            T *So = const_cast<T *>(S) + (size_t)k * 3 * bb + (side ? 2 * bb : 0);
            T *Dn = aug + (side ? 2 : 0) * 2 * bb + bb;
            __syncthreads();
            glass::gemm<T, false>(nx, nx, nx, (T)1, So, Dn, tmp);                       /* S_off * D_n^-1 */
            __syncthreads();
            glass::gemm<T, false>(nx, nx, nx, (T)-1, const_cast<T *>(Dk), tmp, out);    /* -D_k^-1 * (...) */
            __syncthreads();
            for (unsigned i = threadIdx.x; i < bb; i += blockDim.x) P[(size_t)k * 3 * bb + (side ? 2 * bb : 0) + i] = out[i];
        }
    }
}

int ref_glass_build_pinv(const void *d_S, void *d_P, int kind)
{
    const size_t nx = STATE_SIZE;
    const size_t smem = sizeof(T) * (3 * 2 * nx * nx + 2 * nx + 1 + 2 * nx * nx);
    cudaError_t e = cudaFuncSetAttribute(ref_glass_pinv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    ref_glass_pinv_kernel<<<KNOT_POINTS, 64, smem>>>((const T *)d_S, (T *)d_P, kind);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    return (int)e;
}

/* Occupancy facts for the report. */
int ref_kernel_attrs(int *regs, int *max_blocks_per_sm)
{
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, (const void *)pcg<T, STATE_SIZE, KNOT_POINTS>);
    if (e != cudaSuccess) return (int)e;
    *regs = a.numRegs;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(max_blocks_per_sm, (const void *)pcg<T, STATE_SIZE, KNOT_POINTS>,
                                                      pcg_constants::DEFAULT_BLOCK.x, pcgSharedMemSize<T>(STATE_SIZE, KNOT_POINTS));
    return (int)e;
}

}  // extern "C"
