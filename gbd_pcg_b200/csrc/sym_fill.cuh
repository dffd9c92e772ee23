// sym_fill.cuh -- GBDPCG_FLAG_SYMMETRIC_S: the host entry points ship only the diagonal and right blocks of S over PCIe (2/3 of the
// bytes; a strided cudaMemcpy2DAsync) and this kernel rebuilds the left blocks on the device: left(k) = right(k-1)^T, k >= 1.
// PCG needs a symmetric S anyway; the flag is opt-in because the reference kernel itself never assumes it (it reads all three
// blocks of a block-row, include/utils.cuh:46-85) and a caller may hold an S whose stored halves differ in the last bit.
// The padding block left of block-row 0 is zeroed (never read by the solver, include/pcg.cuh:105-106).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gbdpcg {

template <typename T>
__global__ void __launch_bounds__(256) sym_fill_left_kernel(T *__restrict__ S, uint32_t nx, uint32_t N, uint32_t batch)
{
    const uint32_t bb = nx * nx;
    const unsigned long long blocks = (unsigned long long)batch * N;
    for (unsigned long long item = blockIdx.x; item < blocks; item += gridDim.x) {
        const uint32_t k = (uint32_t)(item % N);
        T *row = S + item * 3ull * bb;                 // block-row k of its system: [left | diag | right]
        if (k == 0) {
            for (uint32_t i = threadIdx.x; i < bb; i += blockDim.x) row[i] = (T)0;
        } else {
            const T *right_prev = row - 3ull * bb + 2ull * bb;
            for (uint32_t i = threadIdx.x; i < bb; i += blockDim.x) {
                const uint32_t c = i / nx, r = i - c * nx;         // left(r, c) at c*nx + r  =  right_prev(c, r) at r*nx + c
                row[i] = right_prev[r * nx + c];
            }
        }
    }
}

}  // namespace gbdpcg
