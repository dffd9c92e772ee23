// CSR + device-built preconditioner example:
//   nvcc -std=c++17 -I../include pcg_solve_csr.cu -L../gbd_pcg_b200 -lgbdpcg -o pcg_csr.exe
// Exercises the two host overloads the reference declares but cannot serve: solvePCG(csr_t<T>*, csr_t<T>*, ...)
// (reference include/interface.cuh:8-20: "NOT IMPLEMENTED") and the host-pointer overload with empty_pinv = 0
// (reference include/interface.cuh:33-34: warning + uninitialised Phi^-1).
#include <cmath>
#include <cstdio>
#include <vector>
#include "gpu_pcg.cuh"
#include "example_system.h"

int main()
{
    using namespace example_system;
    typedef double T;
    T S[36], gamma[6], lambda[6];
    fill(S, gamma, lambda);

    // compressed block-tridiagonal -> CSR (row-major walk over the dense 6x6 matrix)
    const unsigned nx = kStateSize, N = kKnotPoints, dim = nx * N;
    std::vector<uint32_t> row_ptr(1, 0), col_ind;
    std::vector<T> val;
    for (unsigned i = 0; i < dim; i++) {
        const unsigned k = i / nx, r = i % nx;
        for (unsigned j = 0; j < dim; j++) {
            const int blk = (int)(j / nx) - (int)k + 1;
            if (blk < 0 || blk > 2) continue;
            const T v = S[k * 3 * nx * nx + blk * nx * nx + (j % nx) * nx + r];
            if (v != 0) { col_ind.push_back(j); val.push_back(v); }
        }
        row_ptr.push_back((uint32_t)val.size());
    }
    csr_t<T> csr = {row_ptr.data(), col_ind.data(), val.data(), dim, dim, (uint32_t)val.size()};

    int rc = 0;
    for (int pass = 0; pass < 2; pass++) {
        pcg_config<T> config;
        config.empty_pinv = 0;                      // no Phi^-1 argument: symmetric-stair Phi^-1 formed on the device
        for (int i = 0; i < 6; i++) lambda[i] = 0;
        const uint32_t iters = pass == 0 ? solvePCG<T>(&csr, (csr_t<T> *)nullptr, gamma, lambda, nx, N, &config)
                                         : solvePCG<T>(S, gamma, lambda, nx, N, &config);
        double num = 0, den = 0;
        for (int i = 0; i < 6; i++) {
            num += (lambda[i] - kLambdaStar[i]) * (lambda[i] - kLambdaStar[i]);
            den += kLambdaStar[i] * kLambdaStar[i];
        }
        const double err = std::sqrt(num / den);
        std::printf("%s: %u iterations, relative error vs S^-1 gamma %.3e\n", pass == 0 ? "CSR overload        " : "host-pointer overload", iters, err);
        if (!(err < 1e-8) || iters != 3) rc = 1;    // symmetric stair, fp64 -> 3 iterations
    }
    return rc;
}
