// types.cuh -- public types of the drop-in headers: csr_t<T> and pcg_config<T>, field for field the
// reference's (include/types.cuh:7-15 and :18-35) so that user code compiles unchanged.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>  // dim3
#include "constants.cuh"
#include "gbdpcg.h"

// Compressed-sparse-row matrix (row_ptr has rows+1 entries).  Accepted by the CSR overload of solvePCG.
template <typename T>
struct csr_t {
    uint32_t *row_ptr;
    uint32_t *col_ind;
    T *val;
    uint32_t rows;
    uint32_t cols;
    uint32_t nnz;
};

// Solver options.  pcg_exit_tol: absolute bound on |r . Phi^-1 r| (strict <); pcg_max_iter: iteration
// cap; pcg_grid / pcg_block: carried, not used (as in the reference); empty_pinv: non-zero = no
// preconditioner supplied -> identity.
// Extension (trailing, defaulted; the reference's five fields keep their names, order and defaults): pinv_build
// selects what the host-pointer overload -- which has no Phi^-1 argument -- does when empty_pinv == 0: it forms
// Phi^-1 from S on the device (GBDPCG_PINV_STAIR by default, or GBDPCG_PINV_JACOBI).  The reference prints a
// warning there and solves with uninitialised memory (include/interface.cuh:33-34,45-46).
template <typename T>
struct pcg_config {
    T pcg_exit_tol;
    uint32_t pcg_max_iter;
    dim3 pcg_grid;
    dim3 pcg_block;
    int empty_pinv;
    int pinv_build = GBDPCG_PINV_STAIR;

    pcg_config(T exit_tol = pcg_constants::DEFAULT_EPSILON<T>, uint32_t max_iter = pcg_constants::DEFAULT_MAX_PCG_ITER,
               dim3 grid = pcg_constants::DEFAULT_GRID, dim3 block = pcg_constants::DEFAULT_BLOCK, int empty_pinv = 1)
        : pcg_exit_tol(exit_tol), pcg_max_iter(max_iter), pcg_grid(grid), pcg_block(block), empty_pinv(empty_pinv)
    {
    }

    // POD form for the C ABI (gbdpcg.h)
    gbdpcg_config to_c() const
    {
        gbdpcg_config c;
        gbdpcg_config_default(&c);
        c.exit_tol = static_cast<double>(pcg_exit_tol);
        c.max_iter = pcg_max_iter;
        c.empty_pinv = empty_pinv;
        return c;
    }
};
