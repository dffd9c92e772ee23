/*
 * gbdpcg.h -- C ABI of the B200-native block-tridiagonal PCG solver (libgbdpcg.so).
 *
 * This is the drop-in boundary for the one hot path of A2R-Lab/GBD-PCG: the solve
 * Phi^-1 S lambda = Phi^-1 gamma.  Plain pointers and sizes only; every entry point names the
 * reference interface it replaces (paths relative to the reference tree).  The C++ header shim
 * (gpu_pcg.cuh / interface.cuh in this directory) forwards the reference's own templates
 * solvePCG<T> / pcg_solve<T> onto these functions; INTEGRATION.md shows the binding.
 *
 * Data layout (reference include/pcg.cuh:104-110, include/utils.cuh:58-84):
 *   S, Pinv : T[3*N*nx*nx]  block-row k at k*3*nx*nx = [left | diag | right], each block column-major;
 *             left block of row 0 and right block of row N-1 are padding and are never read.
 *   gamma, lambda : T[N*nx].  lambda is in/out (warm start, reference include/pcg.cuh:119,215).
 *   batched : problem i lives at S + i*3*N*nx*nx, gamma + i*N*nx, ... (problem-major slabs).
 *
 * Semantics (reference include/pcg.cuh:118-212): absolute exit test |r.Phi^-1 r| < exit_tol (strict),
 * at least one iteration, iters = it+1 on convergence at loop index it, = max_iter otherwise with
 * max_iter_exit = 1; NaN never converges; max_iter = 0 leaves lambda untouched.  Requires N >= 2.
 * Pinv == NULL or cfg->empty_pinv != 0 selects the identity preconditioner (the reference host
 * overload leaves d_Pinv uninitialised, include/interface.cuh:45-46; identity is its evident intent).
 *
 * All functions return GBDPCG_OK (0) or a negative gbdpcg_status; none of them calls exit().
 * There is no CPU fallback: without a usable CUDA device every solve returns GBDPCG_ERR_CUDA.
 */
#ifndef GBDPCG_H
#define GBDPCG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GBDPCG_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define GBDPCG_API __attribute__((visibility("default")))
#else
#define GBDPCG_API
#endif

typedef enum gbdpcg_status {
    GBDPCG_OK = 0,
    GBDPCG_ERR_INVALID = -1,     /* null pointer, nx == 0, N < 2, bad flags                                 */
    GBDPCG_ERR_CUDA = -2,        /* a CUDA runtime call failed; see gbdpcg_last_error() / _last_cuda_error() */
    GBDPCG_ERR_UNSUPPORTED = -3, /* (nx, N, dtype) exceeds what any kernel variant can hold on one GPU       */
    GBDPCG_ERR_NOMEM = -4
} gbdpcg_status;

/* POD mirror of the reference's pcg_config<T> (include/types.cuh:18-35): same meaning and defaults
 * (include/constants.cuh:14-20).  pcg_grid / pcg_block are not carried: the reference ignores them
 * (include/interface.cuh:132 hard-codes the launch shape). */
typedef struct gbdpcg_config {
    double exit_tol;     /* pcg_exit_tol, default 1e-6; compared in T                                   */
    uint32_t max_iter;   /* pcg_max_iter, default 25                                                    */
    int32_t empty_pinv;  /* default 1: ignore Pinv and use identity                                     */
    uint32_t flags;      /* GBDPCG_FLAG_*; 0 = let the library choose                                   */
    uint32_t cluster;    /* 0 = auto; otherwise force thread-block-cluster size 1,2,4,8,16 (tuning/tests) */
} gbdpcg_config;

#define GBDPCG_FLAG_FORCE_GENERIC 0x1u /* use the runtime-nx fallback kernel even if a specialised one exists */
#define GBDPCG_FLAG_PINV_SMEM 0x2u     /* keep Phi^-1 rows in shared memory instead of registers              */
#define GBDPCG_FLAG_PINV_REG 0x4u      /* keep Phi^-1 rows in registers (fail if they do not fit)             */
#define GBDPCG_FLAG_TEXTBOOK 0x8u      /* the reference's order of operations: two reductions per iteration */
#define GBDPCG_FLAG_FUSED 0x10u        /* single-reduction (Chronopoulos-Gear) form of the same recurrence
                                          (the default for both precisions)                                   */
#define GBDPCG_FLAG_LEGACY_ROWS 0x20u  /* first-generation one-row-per-thread kernel (kept for A/B measurements) */
#define GBDPCG_FLAG_NO_TMA 0x40u       /* stage matrices with ordinary loads instead of cp.async.bulk           */
/* Host entry points only: form Phi^-1 from S ON THE DEVICE (gbdpcg_build_pinv_*) instead of taking it from the caller;
 * the Pinv argument and empty_pinv are then ignored.  This is what makes the reference's own host signature
 * solvePCG<T>(h_S, h_gamma, h_lambda, ...) -- which has no Pinv parameter (include/interface.cuh:24-31) -- a
 * PRECONDITIONED solve, and it halves the bytes a host-resident batch has to push over PCIe. */
#define GBDPCG_FLAG_TMEM 0x400u         /* fp32 batched: two (small systems: four) systems per SM with Phi^-1 in tensor memory (pcg_tmem4.cuh: four rows per
                                           thread, S in registers; pcg_tmem.cuh: two rows per thread for systems of more than 128 such threads);
                                           GBDPCG_ERR_UNSUPPORTED for shapes neither covers.  fp64 (even nx 8..16): the tensor-memory form of the cluster
                                           kernel (pcg_fused.cuh, P_TMEM: two rows per thread, Phi^-1 rows in tensor memory, half the SMs per system) --
                                           what the planner itself takes for fp64 batches too large to be resident at once; the same form with four rows
                                           per thread for fp32 systems of even nx 8..16 beyond 256 four-row threads (nx 14: N > 64; C4-sized).  Assumes no other
                                           tensor-memory user (e.g. a concurrent tcgen05 GEMM on another stream) shares the SMs: tcgen05.alloc blocks
                                           until columns are free */
#define GBDPCG_FLAG_TMEM_R2 0x1000u     /* with GBDPCG_FLAG_TMEM or the planner's own choice: the two-rows-per-thread tensor-memory kernel (pcg_tmem.cuh) even where the
                                           four-rows-per-thread one (pcg_tmem4.cuh, systems of up to 128 threads) applies -- A/B measurements */
#define GBDPCG_FLAG_NO_TMEM 0x800u      /* never pick a tensor-memory kernel.  The planner takes one for fp32, Phi^-1 given, even nx <= 16, 16-byte aligned S / Phi^-1:
                                           systems of 3-4 four-row warps (e.g. nx 14, 24 <= N <= 32) at every batch size including the single solve,
                                           other shapes for batches that fill every tensor-memory slot of the GPU */
#define GBDPCG_FLAG_SYMMETRIC_S 0x2000u /* host entry points (_solve_host, _solve_batched_host): the caller declares S symmetric (block (k, k-1) is the
                                           transpose of block (k-1, k), as PCG requires anyway): only the diagonal and right blocks are copied to the
                                           device -- 2/3 of the PCIe bytes -- and the left blocks are rebuilt there by transposition; whatever the
                                           left blocks of h_S hold is ignored.  Opt-in: the reference kernel reads all three blocks and never assumes
                                           symmetry (include/utils.cuh:46-85) */
#define GBDPCG_FLAG_BUILD_JACOBI 0x100u
#define GBDPCG_FLAG_BUILD_STAIR 0x200u

GBDPCG_API void gbdpcg_config_default(gbdpcg_config *cfg);

/* Opaque per-device workspace: cached device buffers, pinned staging, streams.  Replaces the 20
 * allocator calls per solve of the reference host path (include/interface.cuh:39-53,79-86,105-108,140-141).
 * ws == NULL in any call below selects a process-wide default workspace for the current device
 * (internally locked; the reference is not thread-safe at all). */
typedef struct gbdpcg_workspace gbdpcg_workspace;
GBDPCG_API int gbdpcg_workspace_create(gbdpcg_workspace **ws);
GBDPCG_API void gbdpcg_workspace_destroy(gbdpcg_workspace *ws);

/* -- host-pointer solve: replaces solvePCG<T>(T* h_S, T* h_gamma, T* h_lambda, unsigned stateSize,
 *    unsigned knotPoints, pcg_config<T>*)   reference include/interface.cuh:24-89 (README.md:42 pcg_solve<T>).
 *    Blocking.  h_Pinv may be NULL.  Unlike the reference it reports the iteration count (the reference
 *    returns 1 unconditionally, include/interface.cuh:88) and does not copy S / gamma back. */
GBDPCG_API int gbdpcg_solve_host_f32(gbdpcg_workspace *ws, const float *h_S, const float *h_Pinv, const float *h_gamma,
                          float *h_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                          uint32_t *iters, int32_t *max_iter_exit);
GBDPCG_API int gbdpcg_solve_host_f64(gbdpcg_workspace *ws, const double *h_S, const double *h_Pinv, const double *h_gamma,
                          double *h_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                          uint32_t *iters, int32_t *max_iter_exit);

/* -- device-pointer solve: replaces solvePCG<T>(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda,
 *    d_r, d_p, d_v_temp, d_eta_new_temp, config)   reference include/interface.cuh:92-144.
 *    Asynchronous on `stream` (a cudaStream_t; NULL = legacy default stream).  d_iters (uint32) and
 *    d_max_iter_exit (uint8) are device scalars and may be NULL.  The four scratch vectors of the
 *    reference are not needed (r, p and the partial sums never leave the chip). */
GBDPCG_API int gbdpcg_solve_dev_f32(gbdpcg_workspace *ws, const float *d_S, const float *d_Pinv, const float *d_gamma,
                         float *d_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                         uint32_t *d_iters, uint8_t *d_max_iter_exit, void *stream);
GBDPCG_API int gbdpcg_solve_dev_f64(gbdpcg_workspace *ws, const double *d_S, const double *d_Pinv, const double *d_gamma,
                         double *d_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                         uint32_t *d_iters, uint8_t *d_max_iter_exit, void *stream);

/*    Blocking form with the reference's return convention: *iters is what solvePCG returns at
 *    include/interface.cuh:136,143; *max_iter_exit is the flag the reference computes but never reads
 *    back (include/pcg.cuh:212, include/interface.cuh:107,141). */
GBDPCG_API int gbdpcg_solve_dev_sync_f32(gbdpcg_workspace *ws, const float *d_S, const float *d_Pinv, const float *d_gamma,
                              float *d_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                              uint32_t *iters, int32_t *max_iter_exit, void *stream);
GBDPCG_API int gbdpcg_solve_dev_sync_f64(gbdpcg_workspace *ws, const double *d_S, const double *d_Pinv, const double *d_gamma,
                              double *d_lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                              uint32_t *iters, int32_t *max_iter_exit, void *stream);

/* -- batched solves (no reference equivalent: upstream loops over single launches of pcg<>,
 *    include/interface.cuh:132).  `batch` independent systems of identical (nx, N), problem-major slabs.
 *    One CTA (or one thread-block cluster) per problem, persistent over the batch.
 *    _dev: asynchronous on `stream`; d_iters[batch] (uint32) / d_max_iter_exit[batch] (uint8) may be NULL.
 *    _host: blocking; copies are chunked and overlapped with the solves; pointers may be pageable or pinned. */
GBDPCG_API int gbdpcg_solve_batched_dev_f32(gbdpcg_workspace *ws, const float *d_S, const float *d_Pinv, const float *d_gamma,
                                 float *d_lambda, uint32_t nx, uint32_t N, uint32_t batch, const gbdpcg_config *cfg,
                                 uint32_t *d_iters, uint8_t *d_max_iter_exit, void *stream);
GBDPCG_API int gbdpcg_solve_batched_dev_f64(gbdpcg_workspace *ws, const double *d_S, const double *d_Pinv, const double *d_gamma,
                                 double *d_lambda, uint32_t nx, uint32_t N, uint32_t batch, const gbdpcg_config *cfg,
                                 uint32_t *d_iters, uint8_t *d_max_iter_exit, void *stream);
GBDPCG_API int gbdpcg_solve_batched_host_f32(gbdpcg_workspace *ws, const float *h_S, const float *h_Pinv, const float *h_gamma,
                                  float *h_lambda, uint32_t nx, uint32_t N, uint32_t batch, const gbdpcg_config *cfg,
                                  uint32_t *h_iters, uint8_t *h_max_iter_exit);
GBDPCG_API int gbdpcg_solve_batched_host_f64(gbdpcg_workspace *ws, const double *h_S, const double *h_Pinv, const double *h_gamma,
                                  double *h_lambda, uint32_t nx, uint32_t N, uint32_t batch, const gbdpcg_config *cfg,
                                  uint32_t *h_iters, uint8_t *h_max_iter_exit);

/* -- allocation-free device-pointer solve.  NO COUNTERPART IN THE REFERENCE: added for
 *    callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *    One system on the cooperative-grid kernel (pcg_grid.cuh -- the kernel gbdpcg_solve_dev_* takes for systems beyond one 16-CTA
 *    cluster, whose cross-CTA exchange buffers otherwise come from a stream-ordered pool allocation), with those buffers in a
 *    file-scope device array of GBDPCG_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit.  The call
 *    makes no cudaMalloc / cudaFree / pool allocation.  Same arguments and semantics as gbdpcg_solve_dev_f32 / _f64 (is_f64
 *    selects the element type of the four arrays); asynchronous on `stream`; d_iters / d_max_iter_exit may be NULL.
 *    Limits: even nx <= 16, default recurrence (no GBDPCG_FLAG_TEXTBOOK), cfg->cluster == 0, N within the grid kernel's
 *    co-residency limit, and  CTAs * (6*nx + 2) * sizeof(T) <= GBDPCG_STATIC_SCRATCH_BYTES, CTAs = the `cluster` field that
 *    gbdpcg_plan_query reports for a grid-kernel shape (at most one per SM).  On a 148-SM B200 that admits every fp32 shape (at
 *    most 58 KB), every fp64 shape up to nx = 14 (at most 101,824 B) and fp64 nx = 16 on up to 135 CTAs; GBDPCG_ERR_UNSUPPORTED
 *    with a message naming both sizes otherwise (the caller then needs gbdpcg_solve_dev_*).
 *    EVERY call on a device shares the one array (it keeps the last call's contents; the kernel overwrites each word before reading
 *    it, so nothing is cleared): two calls that overlap on the GPU -- different streams, or different host threads -- RACE.
 *    Serialise them on one stream, or use gbdpcg_solve_dev_* for concurrent solves. */
#define GBDPCG_STATIC_SCRATCH_BYTES 106496 /* 104 KiB */
GBDPCG_API int gbdpcg_solve_dev_static_scratch(int is_f64, const void *d_S, const void *d_Pinv, const void *d_gamma, void *d_lambda,
                                               uint32_t nx, uint32_t N, const gbdpcg_config *cfg, uint32_t *d_iters,
                                               uint8_t *d_max_iter_exit, void *stream);

/* -- preconditioner construction on the device (no reference entry point: the reference
 *    carries glass::invertMatrix / glass::gemm / load_block_bd (GLASS/src/L3/inv.cuh:6-37, GLASS/src/L3/gemm.cuh:9-103,
 *    include/utils.cuh:96-161) for exactly this and never calls them).  From S in the compressed layout:
 *      GBDPCG_PINV_IDENTITY  diag blocks = I
 *      GBDPCG_PINV_JACOBI    Phi^-1_kk = S_kk^-1                      (Gauss-Jordan without pivoting, inv.cuh order)
 *      GBDPCG_PINV_STAIR     + Phi^-1_{k,k+-1} = -S_kk^-1 S_{k,k+-1} S_{k+-1,k+-1}^-1   (symmetric stair)
 *    Padding blocks of the output are zero.  `batch` problem-major systems.  _dev is asynchronous on `stream`;
 *    d_Pinv must not alias d_S.  _host is blocking. */
#define GBDPCG_PINV_IDENTITY 0
#define GBDPCG_PINV_JACOBI 1
#define GBDPCG_PINV_STAIR 2
GBDPCG_API int gbdpcg_build_pinv_dev_f32(gbdpcg_workspace *ws, const float *d_S, float *d_Pinv, uint32_t nx, uint32_t N, uint32_t batch,
                                         int kind, void *stream);
GBDPCG_API int gbdpcg_build_pinv_dev_f64(gbdpcg_workspace *ws, const double *d_S, double *d_Pinv, uint32_t nx, uint32_t N, uint32_t batch,
                                         int kind, void *stream);
GBDPCG_API int gbdpcg_build_pinv_host_f32(gbdpcg_workspace *ws, const float *h_S, float *h_Pinv, uint32_t nx, uint32_t N, uint32_t batch,
                                          int kind);
GBDPCG_API int gbdpcg_build_pinv_host_f64(gbdpcg_workspace *ws, const double *h_S, double *h_Pinv, uint32_t nx, uint32_t N, uint32_t batch,
                                          int kind);

/* -- the steps either side of the solve (not in the reference tree: upstream MPCGPU, cited at
 *    reference README.md:66-76, forms S and gamma on the GPU before launching pcg<>, include/interface.cuh:92-103).
 *    KKT system [G C^T; C 0][z; lambda] = [-g; -c] of the LQR sub-problem, G = blkdiag(Q_0, R_0, ..., Q_{N-1}),
 *    C row-block 0 = [I], row-block k+1 = [-A_k -B_k I]:   S = -C G^-1 C^T,  gamma = C G^-1 g - c  (the sign convention
 *    of the reference example, S negative definite):
 *      S_00 = -Q_0^-1,  S_{k+1,k+1} = -(A_k Q_k^-1 A_k^T + B_k R_k^-1 B_k^T + Q_{k+1}^-1),  S_{k+1,k} = A_k Q_k^-1 = S_{k,k+1}^T
 *      gamma_0 = Q_0^-1 q_0 - c_0,  gamma_{k+1} = -A_k Q_k^-1 q_k - B_k R_k^-1 r_k + Q_{k+1}^-1 q_{k+1} - c_{k+1}
 *    _schur_form_dev: stage-major device inputs, every matrix column-major: Q [batch][N][nx*nx], R [batch][N-1][nu*nu],
 *      A [batch][N-1][nx*nx], B [batch][N-1][nx*nu], q, c [batch][N][nx], r [batch][N-1][nu]; writes S (compressed layout,
 *      padding blocks zero) and gamma; d_Qinv [batch][N][nx*nx] / d_Rinv [batch][N-1][nu*nu] are optional by-products.
 *    _schur_recover_dev: x_k = -Q_k^-1 (q_k + lambda_k - A_k^T lambda_{k+1}), u_k = -R_k^-1 (r_k - B_k^T lambda_{k+1});
 *      x [batch][N][nx], u [batch][N-1][nu].   Both asynchronous on `stream`. */
GBDPCG_API int gbdpcg_schur_form_dev_f32(gbdpcg_workspace *ws, const float *d_Q, const float *d_R, const float *d_A, const float *d_B,
                                         const float *d_q, const float *d_r, const float *d_c, uint32_t nx, uint32_t nu, uint32_t N,
                                         uint32_t batch, float *d_S, float *d_gamma, float *d_Qinv, float *d_Rinv, void *stream);
GBDPCG_API int gbdpcg_schur_form_dev_f64(gbdpcg_workspace *ws, const double *d_Q, const double *d_R, const double *d_A, const double *d_B,
                                         const double *d_q, const double *d_r, const double *d_c, uint32_t nx, uint32_t nu, uint32_t N,
                                         uint32_t batch, double *d_S, double *d_gamma, double *d_Qinv, double *d_Rinv, void *stream);
GBDPCG_API int gbdpcg_schur_recover_dev_f32(gbdpcg_workspace *ws, const float *d_Qinv, const float *d_Rinv, const float *d_A,
                                            const float *d_B, const float *d_q, const float *d_r, const float *d_lambda, uint32_t nx,
                                            uint32_t nu, uint32_t N, uint32_t batch, float *d_x, float *d_u, void *stream);
GBDPCG_API int gbdpcg_schur_recover_dev_f64(gbdpcg_workspace *ws, const double *d_Qinv, const double *d_Rinv, const double *d_A,
                                            const double *d_B, const double *d_q, const double *d_r, const double *d_lambda, uint32_t nx,
                                            uint32_t nu, uint32_t N, uint32_t batch, double *d_x, double *d_u, void *stream);

/* -- CSR front end: replaces the reference's unimplemented overload
 *    solvePCG(csr_t<T>* h_S, csr_t<T>* h_Pinv, T* h_gamma, T* h_lambda, stateSize, knotPoints, config)
 *    (include/interface.cuh:8-20 prints "NOT IMPLEMENTED" and exits 12; csr_t<T>: include/types.cuh:7-15).
 *    _pack_csr_dev: device CSR (rows = nx*N, uint32 row_ptr[rows+1] / col_ind[nnz], T val[nnz]) -> compressed
 *      block-tridiagonal layout in d_out (zeroed by the call); duplicates are summed; an entry outside the band sets
 *      bit 0 of *d_status (device int32, may be NULL) and is dropped; a row_ptr that decreases or runs
 *      past `nnz` (the number of entries col_ind / val hold) sets bit 1 and is clamped -- the kernel never reads outside
 *      col_ind[0..nnz) / val[0..nnz).  Asynchronous on `stream`.
 *    _solve_csr_host: blocking solve with S (and optionally Phi^-1) given as host CSR.  Phi^-1 is used when its arrays
 *      are non-NULL and cfg->empty_pinv == 0; GBDPCG_FLAG_BUILD_* forms it from S instead.  Returns
 *      GBDPCG_ERR_INVALID when an entry lies outside the block-tridiagonal band, and -- checked on the host before anything is
 *      uploaded -- when row_ptr[0] != 0, row_ptr decreases, or row_ptr[rows] > nnz. */
GBDPCG_API int gbdpcg_pack_csr_dev_f32(gbdpcg_workspace *ws, const uint32_t *d_row_ptr, const uint32_t *d_col_ind, const float *d_val,
                                       uint32_t nnz, uint32_t rows, uint32_t nx, uint32_t N, float *d_out, int32_t *d_status, void *stream);
GBDPCG_API int gbdpcg_pack_csr_dev_f64(gbdpcg_workspace *ws, const uint32_t *d_row_ptr, const uint32_t *d_col_ind, const double *d_val,
                                       uint32_t nnz, uint32_t rows, uint32_t nx, uint32_t N, double *d_out, int32_t *d_status, void *stream);
GBDPCG_API int gbdpcg_solve_csr_host_f32(gbdpcg_workspace *ws, const uint32_t *S_row_ptr, const uint32_t *S_col_ind, const float *S_val,
                                         uint32_t S_nnz, const uint32_t *P_row_ptr, const uint32_t *P_col_ind, const float *P_val,
                                         uint32_t P_nnz, const float *h_gamma, float *h_lambda, uint32_t nx, uint32_t N,
                                         const gbdpcg_config *cfg, uint32_t *iters, int32_t *max_iter_exit);
GBDPCG_API int gbdpcg_solve_csr_host_f64(gbdpcg_workspace *ws, const uint32_t *S_row_ptr, const uint32_t *S_col_ind, const double *S_val,
                                         uint32_t S_nnz, const uint32_t *P_row_ptr, const uint32_t *P_col_ind, const double *P_val,
                                         uint32_t P_nnz, const double *h_gamma, double *h_lambda, uint32_t nx, uint32_t N,
                                         const gbdpcg_config *cfg, uint32_t *iters, int32_t *max_iter_exit);

/* -- launch-plan query: replaces pcgSharedMemSize<T> (reference include/pcg.cuh:13-20) and
 *    checkPcgOccupancy<T> (include/pcg.cuh:23-49) with what THIS library would launch. */
typedef struct gbdpcg_plan_info {
    int32_t specialised;   /* 1: register-row kernel compiled for this nx; 0: runtime-nx fallback */
    int32_t cluster;       /* CTAs per problem                                                   */
    int32_t threads;       /* threads per CTA                                                    */
    int32_t block_rows;    /* block-rows (knot points) owned by one CTA                          */
    int32_t pinv_mode;     /* 0 identity, 1 registers, 2 shared memory, 3 global (fallback), 4 tensor memory (fp64 throughput form) */
    int32_t regs;          /* registers per thread of the chosen kernel                          */
    uint64_t smem_bytes;   /* dynamic shared memory per CTA                                      */
    int32_t max_resident;  /* problems resident at once on the current device                    */
    char kernel[96];       /* variant name, for logs                                             */
} gbdpcg_plan_info;
GBDPCG_API int gbdpcg_plan_query(uint32_t nx, uint32_t N, int is_f64, int has_pinv, const gbdpcg_config *cfg,
                      gbdpcg_plan_info *out);
/* The same for a batch of `batch` systems.  gbdpcg_plan_query answers for ONE system (what replaces the reference's
 * pcgSharedMemSize / checkPcgOccupancy, include/pcg.cuh:13-49); a system that needs a thread-block cluster gets a
 * larger one when the whole batch is resident in one wave anyway (latency), the smallest that fits otherwise. */
GBDPCG_API int gbdpcg_plan_query_batched(uint32_t nx, uint32_t N, int is_f64, int has_pinv, uint32_t batch,
                              const gbdpcg_config *cfg, gbdpcg_plan_info *out);

/* -- pinned host memory for callers that want full-rate copies through the _host entry points. */
GBDPCG_API void *gbdpcg_host_alloc(size_t bytes);
GBDPCG_API void gbdpcg_host_free(void *p);

/* -- diagnostics */
GBDPCG_API int gbdpcg_version(void);
GBDPCG_API const char *gbdpcg_build_arch(void);       /* "sm_100a" */
GBDPCG_API const char *gbdpcg_last_error(void);       /* thread-local message of the last failing call */
GBDPCG_API int gbdpcg_last_cuda_error(void);          /* thread-local cudaError_t of the last failing call (0 if none) */
GBDPCG_API uint64_t gbdpcg_kernel_launches(void);     /* number of solver kernels this process has launched */

#ifdef __cplusplus
}
#endif
#endif /* GBDPCG_H */
