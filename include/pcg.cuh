// pcg.cuh -- launch-sizing helpers of the drop-in headers.
//
// The reference's include/pcg.cuh holds its kernel plus two host helpers, pcgSharedMemSize<T>
// (:13-20) and checkPcgOccupancy<T> (:23-49).  The kernels of this build live in libgbdpcg.so
// (gbd_pcg_b200/csrc: pcg_fused.cuh for one system per CTA / cluster, pcg_tmem4.cuh / pcg_tmem.cuh for fp32 batches,
// pcg_grid.cuh beyond one cluster); the two helpers keep their names and answer for the kernel this
// library would actually launch, through gbdpcg_plan_query.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include <cuda_runtime.h>
#include "gbdpcg.h"
#include "gpuassert.cuh"
#include "types.cuh"

namespace gbdpcg_shim {
template <typename T>
inline gbdpcg_plan_info plan_or_die(uint32_t state_size, uint32_t knot_points)
{
    static_assert(std::is_same<T, float>::value || std::is_same<T, double>::value, "T must be float or double");
    gbdpcg_plan_info info;
    const int rc = gbdpcg_plan_query(state_size, knot_points, std::is_same<T, double>::value, 1, nullptr, &info);
    if (rc != GBDPCG_OK) {
        // same exit status the reference uses for "does not fit on the device" (include/pcg.cuh:43-46)
        std::printf("[Error] %s\n", gbdpcg_last_error());
        std::exit(6);
    }
    return info;
}
}  // namespace gbdpcg_shim

// Dynamic shared memory per CTA of the solver kernel for this problem shape.
template <typename T>
size_t pcgSharedMemSize(uint32_t state_size, uint32_t knot_points)
{
    return static_cast<size_t>(gbdpcg_shim::plan_or_die<T>(state_size, knot_points).smem_bytes);
}

// True when one system of this shape can be made resident on the current device.  `kernel` and `block`
// are accepted for source compatibility and ignored (the library picks its own kernel and launch shape).
// Exit statuses follow the reference: 5 = device lacks the required launch capability (here: not a
// Blackwell GPU -- the library is built for sm_100a only, compute capability major < 10), 6 = the system does not fit.
template <typename T>
bool checkPcgOccupancy(void * /*kernel*/, dim3 /*block*/, uint32_t state_size, uint32_t knot_points)
{
    int dev = 0, major = 0;
    gpuErrchk(cudaGetDevice(&dev));
    gpuErrchk(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major < 10) {
        std::printf("[Error] Device is not a Blackwell (sm_100a) GPU\n");
        std::exit(5);
    }
    const gbdpcg_plan_info info = gbdpcg_shim::plan_or_die<T>(state_size, knot_points);
    if (info.max_resident < 1) {
        std::printf("Too many knot points ([%u]). Kernel %s cannot become resident.\n", knot_points, info.kernel);
        std::exit(6);
    }
    return true;
}
