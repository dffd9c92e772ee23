// interface.cuh -- host API of the drop-in headers: solvePCG<T> (three overloads, as the reference's
// include/interface.cuh:8-20, :24-89, :92-144) and pcg_solve<T> (the name the reference's README.md:42
// documents).  Every overload is a thin forwarder onto the C ABI of libgbdpcg.so (gbdpcg.h); no CUDA
// runtime call is made from this header except through gpuErrchk in the helpers.
//
// Differences from the reference, all deliberate:
//   * run-time (stateSize, knotPoints) are honoured; -DSTATE_SIZE/-DKNOT_POINTS are not needed;
//   * the host-pointer overload returns the iteration count (the reference returns 1, interface.cuh:88),
//     uses the identity preconditioner when empty_pinv != 0 and, when empty_pinv == 0, a preconditioner it forms
//     from S on the device (config->pinv_build; the reference prints a warning and reads uninitialised device
//     memory, interface.cuh:33-34,45-46), and does not copy S / gamma back;
//   * no allocation per call (cached workspace inside the library);
//   * the CSR overload, a stub in the reference (message + exit(12)), is implemented: CSR -> compressed layout on the device.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <type_traits>
#include <cuda_runtime.h>
#include "gbdpcg.h"
#include "gpuassert.cuh"
#include "pcg.cuh"
#include "types.cuh"

namespace gbdpcg_shim {

// Turn a C-ABI status into the reference's error behaviour: CUDA failures print "GPUassert: ..." and
// exit with the cudaError_t (include/gpuassert.cuh:5-12); anything else prints the message and exits
// with a small positive status (6 = does not fit, as include/pcg.cuh:46; 1 = bad arguments).
inline void die_on_error(int rc, const char *file, int line)
{
    if (rc == GBDPCG_OK) return;
    if (rc == GBDPCG_ERR_CUDA) {
        std::fprintf(stderr, "GPUassert: %s %s %d\n", gbdpcg_last_error(), file, line);
        const int e = gbdpcg_last_cuda_error();
        std::exit(e ? e : 1);
    }
    std::fprintf(stderr, "[Error] %s\n", gbdpcg_last_error());
    std::exit(rc == GBDPCG_ERR_UNSUPPORTED || rc == GBDPCG_ERR_NOMEM ? 6 : 1);
}

template <typename T>
struct abi;
template <>
struct abi<float> {
    static int host(const float *S, const float *P, const float *g, float *l, uint32_t nx, uint32_t N, const gbdpcg_config *c,
                    uint32_t *it, int32_t *ex)
    { return gbdpcg_solve_host_f32(nullptr, S, P, g, l, nx, N, c, it, ex); }
    static int dev(const float *S, const float *P, const float *g, float *l, uint32_t nx, uint32_t N, const gbdpcg_config *c,
                   uint32_t *it, int32_t *ex, void *stream)
    { return gbdpcg_solve_dev_sync_f32(nullptr, S, P, g, l, nx, N, c, it, ex, stream); }
    static int csr(const uint32_t *srp, const uint32_t *sci, const float *sv, uint32_t snnz, const uint32_t *prp, const uint32_t *pci,
                   const float *pv, uint32_t pnnz, const float *g, float *l, uint32_t nx, uint32_t N, const gbdpcg_config *c, uint32_t *it,
                   int32_t *ex)
    { return gbdpcg_solve_csr_host_f32(nullptr, srp, sci, sv, snnz, prp, pci, pv, pnnz, g, l, nx, N, c, it, ex); }
};
template <>
struct abi<double> {
    static int host(const double *S, const double *P, const double *g, double *l, uint32_t nx, uint32_t N, const gbdpcg_config *c,
                    uint32_t *it, int32_t *ex)
    { return gbdpcg_solve_host_f64(nullptr, S, P, g, l, nx, N, c, it, ex); }
    static int dev(const double *S, const double *P, const double *g, double *l, uint32_t nx, uint32_t N, const gbdpcg_config *c,
                   uint32_t *it, int32_t *ex, void *stream)
    { return gbdpcg_solve_dev_sync_f64(nullptr, S, P, g, l, nx, N, c, it, ex, stream); }
    static int csr(const uint32_t *srp, const uint32_t *sci, const double *sv, uint32_t snnz, const uint32_t *prp, const uint32_t *pci,
                   const double *pv, uint32_t pnnz, const double *g, double *l, uint32_t nx, uint32_t N, const gbdpcg_config *c,
                   uint32_t *it, int32_t *ex)
    { return gbdpcg_solve_csr_host_f64(nullptr, srp, sci, sv, snnz, prp, pci, pv, pnnz, g, l, nx, N, c, it, ex); }
};

}  // namespace gbdpcg_shim

// ---- overload 1: S and Phi^-1 in CSR form (reference include/interface.cuh:8-20: a stub that prints
//      "NOT IMPLEMENTED" and exits 12).  Implemented here: the CSR arrays (host pointers inside csr_t<T>,
//      include/types.cuh:7-15) are packed into the compressed block-tridiagonal layout on the device.
//      h_Pinv may be NULL; it is applied when config->empty_pinv == 0.  With h_Pinv == NULL and
//      config->empty_pinv == 0 the library forms config->pinv_build from S on the device.
template <typename T>
uint32_t solvePCG(csr_t<T> *h_S, csr_t<T> *h_Pinv, T *h_gamma, T *h_lambda, unsigned stateSize, unsigned knotPoints,
                  pcg_config<T> *config)
{
    pcg_config<T> defaults;
    if (!config) config = &defaults;
    if (!h_S) { std::fprintf(stderr, "[Error] h_S must not be NULL\n"); std::exit(1); }
    gbdpcg_config c = config->to_c();
    if (!h_Pinv) {
        if (!config->empty_pinv && config->pinv_build != GBDPCG_PINV_IDENTITY)
            c.flags |= (config->pinv_build == GBDPCG_PINV_JACOBI) ? GBDPCG_FLAG_BUILD_JACOBI : GBDPCG_FLAG_BUILD_STAIR;
        else
            c.empty_pinv = 1;
    }
    if (h_S->rows != stateSize * knotPoints || h_S->cols != h_S->rows) {
        std::fprintf(stderr, "[Error] CSR S must be square with stateSize*knotPoints rows\n");
        std::exit(1);
    }
    uint32_t iters = 0;
    int32_t hit_max = 0;
    gbdpcg_shim::die_on_error(gbdpcg_shim::abi<T>::csr(h_S->row_ptr, h_S->col_ind, h_S->val, h_S->nnz, h_Pinv ? h_Pinv->row_ptr : nullptr,
                                                       h_Pinv ? h_Pinv->col_ind : nullptr, h_Pinv ? h_Pinv->val : nullptr,
                                                       h_Pinv ? h_Pinv->nnz : 0u, h_gamma, h_lambda, stateSize, knotPoints, &c, &iters,
                                                       &hit_max),
                              __FILE__, __LINE__);
    return iters;
}

// ---- overload 2: host pointers, S in the compressed block-tridiagonal layout -----------------------
// h_lambda is in/out (warm start).  Blocking.  Returns the number of PCG iterations executed.
template <typename T>
uint32_t solvePCG(T *h_S, T *h_gamma, T *h_lambda, unsigned stateSize, unsigned knotPoints, pcg_config<T> *config)
{
    pcg_config<T> defaults;
    if (!config) config = &defaults;
    gbdpcg_config c = config->to_c();
    if (!config->empty_pinv && config->pinv_build != GBDPCG_PINV_IDENTITY)
        // no Phi^-1 argument on this overload: the library forms it from S on the device (gbdpcg_build_pinv_*)
        c.flags |= (config->pinv_build == GBDPCG_PINV_JACOBI) ? GBDPCG_FLAG_BUILD_JACOBI : GBDPCG_FLAG_BUILD_STAIR;
    else
        c.empty_pinv = 1;  // identity
    uint32_t iters = 0;
    int32_t hit_max = 0;
    gbdpcg_shim::die_on_error(gbdpcg_shim::abi<T>::host(h_S, nullptr, h_gamma, h_lambda, stateSize, knotPoints, &c, &iters, &hit_max),
                              __FILE__, __LINE__);
    return iters;
}

// Extension of overload 2 (not in the reference): the caller supplies Phi^-1 on the host, same layout as S.
// It is applied when config->empty_pinv == 0.  `max_iter_exit` (nullable) receives the flag the reference
// computes on the device but never returns (include/pcg.cuh:212).
template <typename T>
uint32_t solvePCG(T *h_S, T *h_Pinv, T *h_gamma, T *h_lambda, unsigned stateSize, unsigned knotPoints, pcg_config<T> *config,
                  bool *max_iter_exit = nullptr)
{
    pcg_config<T> defaults;
    if (!config) config = &defaults;
    const gbdpcg_config c = config->to_c();
    uint32_t iters = 0;
    int32_t hit_max = 0;
    gbdpcg_shim::die_on_error(gbdpcg_shim::abi<T>::host(h_S, h_Pinv, h_gamma, h_lambda, stateSize, knotPoints, &c, &iters, &hit_max),
                              __FILE__, __LINE__);
    if (max_iter_exit) *max_iter_exit = hit_max != 0;
    return iters;
}

// ---- overload 3: device pointers ---------------------------------------------------------------------
// Blocking, legacy default stream, returns the iteration count -- as the reference (interface.cuh:92-144).
// d_Pinv is always applied (the reference kernel has no notion of empty_pinv: the flag is passed as a 13th
// launch argument to a 12-parameter kernel, interface.cuh:126 vs pcg.cuh:56-68).  The four scratch buffers
// are accepted and left untouched: r, p and the partial sums never leave the chip in this implementation.
template <typename T>
uint32_t solvePCG(const uint32_t state_size, const uint32_t knot_points, T *d_S, T *d_Pinv, T *d_gamma, T *d_lambda, T * /*d_r*/,
                  T * /*d_p*/, T * /*d_v_temp*/, T * /*d_eta_new_temp*/, pcg_config<T> *config)
{
    pcg_config<T> defaults;
    if (!config) config = &defaults;
    gbdpcg_config c = config->to_c();
    c.empty_pinv = (d_Pinv == nullptr);
    uint32_t iters = 0;
    int32_t hit_max = 0;
    gbdpcg_shim::die_on_error(
        gbdpcg_shim::abi<T>::dev(d_S, d_Pinv, d_gamma, d_lambda, state_size, knot_points, &c, &iters, &hit_max, nullptr), __FILE__,
        __LINE__);
    return iters;
}

// ---- README spelling (reference README.md:42): int pcg_solve<T>(h_S, h_gamma, h_lambda, stateSize,
//      knotPoints, config = default) --------------------------------------------------------------------
template <typename T>
int pcg_solve(T *h_S, T *h_gamma, T *h_lambda, unsigned stateSize, unsigned knotPoints, pcg_config<T> *config = nullptr)
{
    return static_cast<int>(solvePCG<T>(h_S, h_gamma, h_lambda, stateSize, knotPoints, config));
}
