// pinv_build.h -- host entry of the preconditioner builder (pinv_build.cu); called by capi.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gbdpcg {

struct PinvLaunchInfo {
    int specialised, threads, chunk, grid;
    size_t smem;
    const char *name;
};

// Returns 0 on success; a cudaError_t (> 0) when a CUDA call failed; -3 when no kernel can hold the system.
// `launched` (if not NULL) is incremented per kernel launch.  info may be NULL.
template <typename T>
int launch_pinv_build(const T *d_S, T *d_Pinv, uint32_t nx, uint32_t N, uint32_t batch, int kind, cudaStream_t stream,
                      PinvLaunchInfo *info, bool plan_only);

}  // namespace gbdpcg
