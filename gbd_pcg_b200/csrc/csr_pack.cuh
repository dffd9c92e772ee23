// csr_pack.cuh -- CSR -> compressed block-tridiagonal layout on the device.
//
// Front end for the overload the reference declares and never implemented:
//   solvePCG(csr_t<T>* h_S, csr_t<T>* h_Pinv, T* h_gamma, T* h_lambda, stateSize, knotPoints, config)
//   reference include/interface.cuh:8-20 ("NOT IMPLEMENTED", exit(12)), csr_t<T> at include/types.cuh:7-15.
// Target layout: reference include/pcg.cuh:104-110 / include/utils.cuh:58-84 -- block-row k at k*3*nx*nx =
// [left | diag | right], each block column-major.  One warp per CSR row; an entry (i, j) goes to block-row i/nx, block
// (j/nx - i/nx + 1), element (j%nx)*nx + i%nx.  Entries outside the block-tridiagonal band (or outside the matrix) raise
// bit 0 of *status and are dropped; a row_ptr that decreases or runs past nnz raises bit 1 and is clamped (no read outside
// col_ind[0..nnz) / val[0..nnz)); duplicate entries are summed (the usual CSR convention), so the output buffer
// must be zeroed before the launch (the C ABI does it).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gbdpcg {

template <typename T>
__global__ void __launch_bounds__(256) csr_pack_kernel(const uint32_t *__restrict__ row_ptr, const uint32_t *__restrict__ col_ind,
                                                       const T *__restrict__ val, uint32_t nnz, uint32_t rows, uint32_t nx, uint32_t N,
                                                       T *__restrict__ out, int32_t *__restrict__ status)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t dim = nx * N;
    const size_t bb = (size_t)nx * nx;
    bool bad = false, bad_ptr = false;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < rows; i += warps) {
        uint32_t lo = row_ptr[i], hi = row_ptr[i + 1];
        // a malformed row_ptr (decreasing, or running past the nnz entries the caller allocated) must not become an
        // out-of-bounds read of col_ind / val: clamp to [0, nnz] and report through bit 1 of *status
        if (lo > hi || hi > nnz) { bad_ptr = true; hi = hi > nnz ? nnz : hi; lo = lo > hi ? hi : lo; }
        const uint32_t k = i / nx, r = i - k * nx;
        for (uint32_t e = lo + lane; e < hi; e += 32) {
            const uint32_t j = col_ind[e];
            const uint32_t kc = j / nx, c = j - kc * nx;
            if (i >= dim || j >= dim || kc + 1 < k || kc > k + 1) { bad = true; continue; }
            atomicAdd(out + (size_t)k * 3 * bb + (size_t)(kc + 1 - k) * bb + (size_t)c * nx + r, val[e]);
        }
    }
    if (bad && status) atomicOr(status, 1);
    if (bad_ptr && status) atomicOr(status, 2);
}

}  // namespace gbdpcg
