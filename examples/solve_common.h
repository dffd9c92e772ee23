// Shared body of examples/pcg_solve.cu (float) and examples/pcg_solve_dp.cu (double): one solve through the
// drop-in host API, checked against the known solution.
#pragma once
#include <cmath>
#include <cstdio>
#include "gpu_pcg.cuh"
#include "example_system.h"

template <typename T>
int run_example(double accept_rel_err)
{
    using namespace example_system;
    T S[36], gamma[6], lambda[6];
    fill(S, gamma, lambda);

    pcg_config<T> config;                       // defaults: tol 1e-6, 25 iterations, no preconditioner
    const uint32_t iters = solvePCG<T>(S, gamma, lambda, kStateSize, kKnotPoints, &config);

    double num = 0, den = 0;
    for (int i = 0; i < 6; i++) {
        num += (lambda[i] - kLambdaStar[i]) * (lambda[i] - kLambdaStar[i]);
        den += kLambdaStar[i] * kLambdaStar[i];
    }
    const double err = std::sqrt(num / den);
    std::printf("solvePCG<%s>: %u iterations, lambda =", sizeof(T) == 8 ? "double" : "float", iters);
    for (int i = 0; i < 6; i++) std::printf(" %.6f", (double)lambda[i]);
    std::printf("\nrelative error vs S^-1 gamma: %.3e (%s)\n", err, err < accept_rel_err ? "ok" : "TOO LARGE");
    return err < accept_rel_err ? 0 : 1;
}
