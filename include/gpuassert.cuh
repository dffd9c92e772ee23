// gpuassert.cuh -- error convention of the drop-in headers.
// Same contract as the reference's include/gpuassert.cuh:5-14: a failing CUDA call prints
// "GPUassert: <message> <file> <line>" to stderr and terminates the process with the cudaError_t as
// exit status; gpuErrchk(x) wraps a call.  (The C ABI underneath, gbdpcg.h, never exits: it returns status codes.)
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

inline void gpuAssert(cudaError_t code, const char *file, int line, bool abort = true)
{
    if (code == cudaSuccess) return;
    std::fprintf(stderr, "GPUassert: %s %s %d\n", cudaGetErrorString(code), file, line);
    if (abort) std::exit(static_cast<int>(code));
}

#define gpuErrchk(ans) do { gpuAssert((ans), __FILE__, __LINE__); } while (0)
