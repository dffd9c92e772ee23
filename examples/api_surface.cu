// api_surface.cu -- every template of the drop-in headers instantiated and RUN, each result checked against a known answer:
//   1. solvePCG<T> overload 3, device pointers with real d_r / d_p / d_v_temp / d_eta_new_temp scratch
//      (reference include/interface.cuh:92-103 -- the seam upstream MPCGPU calls), float and double;
//   2. pcg_solve<T> with the config defaulted (reference README.md:42), float and double;
//   3. pcgSharedMemSize<T> and checkPcgOccupancy<T> (reference include/pcg.cuh:13-49);
//   4. two host threads, each with its own gbdpcg_workspace and its own stream, solving DIFFERENT systems concurrently
//      through the host-pointer and the device-pointer C ABI (the reference is not thread-safe at all: legacy default
//      stream, header-defined mutable globals);
//   5. --latency FILE nx N tol max_iter reps : p50 / p99 of the blocking device-pointer call (overload 3) on the system in
//      FILE (S, Pinv, gamma as raw float32, written by bench.py), measured here in C++ -- no ctypes in the path.
// Exit status 0 = every check passed.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "example_system.h"
#include "gpu_pcg.cuh"

static int g_fail = 0;
#define CHECK(cond, ...)                                                              \
    do {                                                                              \
        if (!(cond)) { std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); g_fail++; } \
    } while (0)

// A block-tridiagonal negative-definite system with a KNOWN solution: random diagonally dominant blocks, symmetric couplings,
// lambda* drawn at random, gamma = S lambda* formed in double.  Compressed layout of reference include/pcg.cuh:104-110.
template <typename T>
struct KnownSystem {
    unsigned nx, N;
    std::vector<T> S, Pinv, gamma;
    std::vector<double> lambda_star;
};
template <typename T>
static KnownSystem<T> make_known(unsigned nx, unsigned N, unsigned seed)
{
    KnownSystem<T> k;
    k.nx = nx; k.N = N;
    const size_t bs = (size_t)nx * nx;
    std::vector<double> S(3 * N * bs, 0.0), P(3 * N * bs, 0.0);
    std::mt19937 rng(seed);
    std::uniform_real_distribution<double> u(-1.0, 1.0);
    auto at = [&](std::vector<double> &M, unsigned b, unsigned j, unsigned r, unsigned c) -> double & { return M[(size_t)b * 3 * bs + j * bs + (size_t)c * nx + r]; };
    for (unsigned b = 0; b < N; b++) {
        for (unsigned r = 0; r < nx; r++)
            for (unsigned c = r; c < nx; c++) {
                const double v = (r == c) ? -(4.0 + nx + u(rng)) : 0.5 * u(rng);
                at(S, b, 1, r, c) = v;
                at(S, b, 1, c, r) = v;
            }
        if (b + 1 < N)
            for (unsigned r = 0; r < nx; r++)
                for (unsigned c = 0; c < nx; c++) {
                    const double v = 0.7 * u(rng);
                    at(S, b, 2, r, c) = v;          // block (b, b+1)
                    at(S, b + 1, 0, c, r) = v;      // block (b+1, b) = transpose
                }
        for (unsigned r = 0; r < nx; r++) at(P, b, 1, r, r) = 1.0 / at(S, b, 1, r, r);   // point-Jacobi preconditioner
    }
    k.lambda_star.resize((size_t)nx * N);
    for (auto &v : k.lambda_star) v = u(rng);
    std::vector<double> g((size_t)nx * N, 0.0);
    for (unsigned b = 0; b < N; b++)
        for (unsigned j = 0; j < 3; j++) {
            if ((b == 0 && j == 0) || (b == N - 1 && j == 2)) continue;
            for (unsigned c = 0; c < nx; c++)
                for (unsigned r = 0; r < nx; r++) g[(size_t)b * nx + r] += at(S, b, j, r, c) * k.lambda_star[(size_t)(b + j - 1) * nx + c];
        }
    k.S.assign(S.begin(), S.end());
    k.Pinv.assign(P.begin(), P.end());
    k.gamma.assign(g.begin(), g.end());
    return k;
}
template <typename T>
static double rel_err(const T *x, const std::vector<double> &ref)
{
    double num = 0, den = 0;
    for (size_t i = 0; i < ref.size(); i++) { num += ((double)x[i] - ref[i]) * ((double)x[i] - ref[i]); den += ref[i] * ref[i]; }
    return std::sqrt(num / den);
}
template <typename T>
static T *to_device(const std::vector<T> &h)
{
    T *d = nullptr;
    gpuErrchk(cudaMalloc(&d, h.size() * sizeof(T)));
    gpuErrchk(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

// ---- 1. overload 3 with the caller's scratch buffers (reference include/interface.cuh:92-103) ----
template <typename T>
static void check_device_overload(unsigned nx, unsigned N, double tol, double accept)
{
    const KnownSystem<T> k = make_known<T>(nx, N, 7 + nx + N);
    T *d_S = to_device(k.S), *d_P = to_device(k.Pinv), *d_g = to_device(k.gamma);
    T *d_l, *d_r, *d_p, *d_v, *d_e;
    const size_t vb = (size_t)nx * N * sizeof(T);
    gpuErrchk(cudaMalloc(&d_l, vb)); gpuErrchk(cudaMalloc(&d_r, vb)); gpuErrchk(cudaMalloc(&d_p, vb));
    gpuErrchk(cudaMalloc(&d_v, std::max(N, nx) * sizeof(T))); gpuErrchk(cudaMalloc(&d_e, std::max(N, nx) * sizeof(T)));
    gpuErrchk(cudaMemset(d_l, 0, vb));
    gpuErrchk(cudaMemset(d_r, 0x5a, vb)); gpuErrchk(cudaMemset(d_p, 0x5a, vb));   // scratch is the caller's: it must come back untouched or at least harmlessly
    pcg_config<T> config;
    config.pcg_exit_tol = (T)tol;
    config.pcg_max_iter = 500;
    const uint32_t iters = solvePCG<T>(nx, N, d_S, d_P, d_g, d_l, d_r, d_p, d_v, d_e, &config);
    std::vector<T> lam((size_t)nx * N);
    gpuErrchk(cudaMemcpy(lam.data(), d_l, vb, cudaMemcpyDeviceToHost));
    const double err = rel_err(lam.data(), k.lambda_star);
    std::printf("overload 3 <%s> nx=%u N=%u: %u iterations, rel err vs known lambda* %.2e\n", sizeof(T) == 8 ? "double" : "float", nx, N, iters, err);
    CHECK(iters > 1 && iters < 500, "iteration count %u", iters);
    CHECK(err < accept, "rel err %.3e", err);
    // warm start from the solution: the exit test must fire after the first iteration (reference pcg.cuh:195 tests after the update)
    const uint32_t again = solvePCG<T>(nx, N, d_S, d_P, d_g, d_l, d_r, d_p, d_v, d_e, &config);
    CHECK(again <= 2, "warm start took %u iterations", again);
    // d_Pinv == nullptr: identity preconditioner, more iterations, same answer
    gpuErrchk(cudaMemset(d_l, 0, vb));
    const uint32_t it_id = solvePCG<T>(nx, N, d_S, (T *)nullptr, d_g, d_l, d_r, d_p, d_v, d_e, &config);
    gpuErrchk(cudaMemcpy(lam.data(), d_l, vb, cudaMemcpyDeviceToHost));
    CHECK(rel_err(lam.data(), k.lambda_star) < accept, "identity preconditioner rel err %.3e", rel_err(lam.data(), k.lambda_star));
    CHECK(it_id + 1 >= iters && it_id < 500, "identity %u vs Jacobi %u iterations", it_id, iters);
    for (T *p : {d_S, d_P, d_g, d_l, d_r, d_p, d_v, d_e}) cudaFree(p);
}

// ---- 2. pcg_solve<T> with the config defaulted (reference README.md:42) ----
template <typename T>
static void check_readme_name(double accept)
{
    using namespace example_system;
    T S[36], gamma[6], lambda[6];
    fill(S, gamma, lambda);
    const int iters = pcg_solve<T>(S, gamma, lambda, kStateSize, kKnotPoints);     // default config: tol 1e-6, 25 iterations, identity
    std::vector<double> star(kLambdaStar, kLambdaStar + 6);
    const double err = rel_err(lambda, star);
    std::printf("pcg_solve<%s>(default config): %d iterations, rel err %.2e\n", sizeof(T) == 8 ? "double" : "float", iters, err);
    CHECK(iters >= 3 && iters <= 25, "iterations %d", iters);
    CHECK(err < accept, "rel err %.3e", err);
}

// ---- 3. pcgSharedMemSize<T> / checkPcgOccupancy<T> (reference include/pcg.cuh:13-49) ----
static void check_sizing_helpers()
{
    const size_t s32 = pcgSharedMemSize<float>(14, 32), s64 = pcgSharedMemSize<double>(14, 64);
    std::printf("pcgSharedMemSize<float>(14,32) = %zu B, <double>(14,64) = %zu B\n", s32, s64);
    CHECK(s32 > 0 && s32 <= 227 * 1024, "fp32 smem %zu", s32);
    CHECK(s64 > 0 && s64 <= 227 * 1024, "fp64 smem %zu", s64);
    CHECK(checkPcgOccupancy<float>(nullptr, dim3(64), 14, 32), "C2 must fit");
    CHECK(checkPcgOccupancy<double>(nullptr, dim3(64), 14, 64), "C3 must fit");
    CHECK(checkPcgOccupancy<float>(nullptr, dim3(64), 12, 512), "C4 must fit");
}

// ---- 4. two host threads, own workspace + own stream each, different systems, concurrently ----
struct Worker {
    unsigned nx, N, seed;
    int fails = 0;
    double worst = 0;
    unsigned solves = 0;
};
static void worker_main(Worker *w)
{
    gbdpcg_workspace *ws = nullptr;
    if (gbdpcg_workspace_create(&ws) != GBDPCG_OK) { w->fails++; return; }
    cudaStream_t st;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { w->fails++; return; }
    gbdpcg_config cfg;
    gbdpcg_config_default(&cfg);
    cfg.exit_tol = 1e-9; cfg.max_iter = 500; cfg.empty_pinv = 0;
    for (int rep = 0; rep < 40; rep++) {
        const KnownSystem<float> k = make_known<float>(w->nx, w->N, w->seed + rep);
        // (a) host pointers through this thread's workspace
        std::vector<float> lam((size_t)w->nx * w->N, 0.f);
        uint32_t it = 0; int32_t ex = 1;
        int rc = gbdpcg_solve_host_f32(ws, k.S.data(), k.Pinv.data(), k.gamma.data(), lam.data(), w->nx, w->N, &cfg, &it, &ex);
        double e = rel_err(lam.data(), k.lambda_star);
        if (rc != GBDPCG_OK || ex || e > 2e-5) w->fails++;
        w->worst = std::max(w->worst, e);
        // (b) device pointers, blocking form, this thread's stream
        float *d_S = to_device(k.S), *d_P = to_device(k.Pinv), *d_g = to_device(k.gamma), *d_l;
        cudaMalloc(&d_l, lam.size() * sizeof(float));
        cudaMemsetAsync(d_l, 0, lam.size() * sizeof(float), st);
        rc = gbdpcg_solve_dev_sync_f32(ws, d_S, d_P, d_g, d_l, w->nx, w->N, &cfg, &it, &ex, st);
        cudaMemcpyAsync(lam.data(), d_l, lam.size() * sizeof(float), cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        e = rel_err(lam.data(), k.lambda_star);
        if (rc != GBDPCG_OK || ex || e > 2e-5) w->fails++;
        w->worst = std::max(w->worst, e);
        w->solves += 2;
        for (float *p : {d_S, d_P, d_g, d_l}) cudaFree(p);
    }
    cudaStreamDestroy(st);
    gbdpcg_workspace_destroy(ws);
}
static void check_concurrent_workspaces()
{
    Worker a{14, 32, 1000}, b{12, 40, 5000};
    std::thread ta(worker_main, &a), tb(worker_main, &b);
    ta.join(); tb.join();
    std::printf("two threads, own workspace + stream each: %u + %u solves, failures %d + %d, worst rel err %.2e / %.2e\n", a.solves, b.solves,
                a.fails, b.fails, a.worst, b.worst);
    CHECK(a.fails == 0 && b.fails == 0 && a.solves == 80 && b.solves == 80, "concurrent solves failed");
}

// ---- 5. latency of the blocking device-pointer API, measured in C++ ----
static int latency_mode(const char *path, unsigned nx, unsigned N, double tol, unsigned max_iter, int reps)
{
    const size_t me = 3 * (size_t)N * nx * nx, ve = (size_t)N * nx;
    std::vector<float> S(me), P(me), g(ve);
    FILE *f = std::fopen(path, "rb");
    if (!f || std::fread(S.data(), 4, me, f) != me || std::fread(P.data(), 4, me, f) != me || std::fread(g.data(), 4, ve, f) != ve) {
        std::printf("cannot read %s\n", path);
        return 2;
    }
    std::fclose(f);
    float *d_S = to_device(S), *d_P = to_device(P), *d_g = to_device(g), *d_l, *d_scratch;
    gpuErrchk(cudaMalloc(&d_l, ve * 4)); gpuErrchk(cudaMalloc(&d_scratch, ve * 4));
    pcg_config<float> config;
    config.pcg_exit_tol = (float)tol;
    config.pcg_max_iter = max_iter;
    gbdpcg_workspace *ws = nullptr;
    gbdpcg_workspace_create(&ws);
    cudaStream_t st;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    gbdpcg_config c = config.to_c();
    c.empty_pinv = 0;
    std::vector<double> t_shim, t_ws;
    uint32_t iters = 0, it2 = 0;
    int32_t ex = 0;
    for (int i = 0; i < reps + 50; i++) {
        gpuErrchk(cudaMemset(d_l, 0, ve * 4));
        gpuErrchk(cudaDeviceSynchronize());
        auto t0 = std::chrono::steady_clock::now();
        iters = solvePCG<float>(nx, N, d_S, d_P, d_g, d_l, d_scratch, d_scratch, d_scratch, d_scratch, &config);
        auto t1 = std::chrono::steady_clock::now();
        if (i >= 50) t_shim.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
        gpuErrchk(cudaMemsetAsync(d_l, 0, ve * 4, st));
        gpuErrchk(cudaStreamSynchronize(st));
        t0 = std::chrono::steady_clock::now();
        gbdpcg_solve_dev_sync_f32(ws, d_S, d_P, d_g, d_l, nx, N, &c, &it2, &ex, st);
        t1 = std::chrono::steady_clock::now();
        if (i >= 50) t_ws.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
    }
    auto pct = [](std::vector<double> &v, double q) { std::sort(v.begin(), v.end()); return v[(size_t)(q * (v.size() - 1))]; };
    std::printf("{\"iters\": %u, \"iters_ws\": %u, \"shim_overload3_p50_us\": %.2f, \"shim_overload3_p99_us\": %.2f, \"c_abi_ws_stream_p50_us\": %.2f, "
                "\"c_abi_ws_stream_p99_us\": %.2f, \"reps\": %d}\n",
                iters, it2, pct(t_shim, 0.5), pct(t_shim, 0.99), pct(t_ws, 0.5), pct(t_ws, 0.99), reps);
    return 0;
}

// ---- 5. the allocation-free entry point (gbdpcg_solve_dev_static_scratch: no counterpart in the reference), called from C++: a plain call
//         and -- what it is for -- a call recorded under CUDA stream capture and replayed as a graph ----
static void check_static_scratch(unsigned nx, unsigned N, double accept)
{
    const KnownSystem<float> k = make_known<float>(nx, N, 99);
    float *d_S = to_device(k.S), *d_P = to_device(k.Pinv), *d_g = to_device(k.gamma), *d_l;
    uint32_t *d_it;
    const size_t vb = (size_t)nx * N * sizeof(float);
    gpuErrchk(cudaMalloc(&d_l, vb)); gpuErrchk(cudaMalloc(&d_it, sizeof(uint32_t)));
    gbdpcg_config c;
    gbdpcg_config_default(&c);
    c.exit_tol = 1e-9; c.max_iter = 500; c.empty_pinv = 0;
    cudaStream_t st;
    gpuErrchk(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    gpuErrchk(cudaMemsetAsync(d_l, 0, vb, st));
    int rc = gbdpcg_solve_dev_static_scratch(0, d_S, d_P, d_g, d_l, nx, N, &c, d_it, nullptr, st);   // also the per-device set-up call
    CHECK(rc == GBDPCG_OK, "static scratch: %s", gbdpcg_last_error());
    gpuErrchk(cudaStreamSynchronize(st));
    std::vector<float> lam((size_t)nx * N);
    uint32_t it1 = 0;
    gpuErrchk(cudaMemcpy(lam.data(), d_l, vb, cudaMemcpyDeviceToHost)); gpuErrchk(cudaMemcpy(&it1, d_it, 4, cudaMemcpyDeviceToHost));
    const double e1 = rel_err(lam.data(), k.lambda_star);
    CHECK(it1 > 1 && it1 < 500 && e1 < accept, "static scratch plain call: %u iterations, rel err %.3e", it1, e1);
    // under capture: nothing in the call may allocate (an allocating call would fail the capture)
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    gpuErrchk(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    gpuErrchk(cudaMemsetAsync(d_l, 0, vb, st));
    rc = gbdpcg_solve_dev_static_scratch(0, d_S, d_P, d_g, d_l, nx, N, &c, d_it, nullptr, st);
    const cudaError_t ce = cudaStreamEndCapture(st, &graph);
    CHECK(rc == GBDPCG_OK && ce == cudaSuccess && graph != nullptr, "capture: rc %d (%s), cudaStreamEndCapture %d", rc, gbdpcg_last_error(), (int)ce);
    if (graph) {
        gpuErrchk(cudaGraphInstantiate(&exec, graph, 0));
        for (int rep = 0; rep < 3; rep++) gpuErrchk(cudaGraphLaunch(exec, st));
        gpuErrchk(cudaStreamSynchronize(st));
        std::vector<float> lam2((size_t)nx * N);
        uint32_t it2 = 0;
        gpuErrchk(cudaMemcpy(lam2.data(), d_l, vb, cudaMemcpyDeviceToHost)); gpuErrchk(cudaMemcpy(&it2, d_it, 4, cudaMemcpyDeviceToHost));
        CHECK(it2 == it1 && std::memcmp(lam2.data(), lam.data(), vb) == 0, "graph replay differs from the plain call (%u vs %u iterations)", it2, it1);
        cudaGraphExecDestroy(exec);
        cudaGraphDestroy(graph);
    }
    std::printf("static scratch nx=%u N=%u: %u iterations, rel err %.2e, replayed 3x from a captured graph: identical\n", nx, N, it1, e1);
    // beyond the static array / the grid kernel: refused, nothing launched
    rc = gbdpcg_solve_dev_static_scratch(0, d_S, d_P, d_g, d_l, 3, 8, &c, nullptr, nullptr, st);
    CHECK(rc == GBDPCG_ERR_UNSUPPORTED, "odd state size must be refused (rc %d)", rc);
    cudaStreamDestroy(st);
    for (void *p : {(void *)d_S, (void *)d_P, (void *)d_g, (void *)d_l, (void *)d_it}) cudaFree(p);
}

int main(int argc, char **argv)
{
    if (argc >= 8 && std::string(argv[1]) == "--latency")
        return latency_mode(argv[2], (unsigned)std::atoi(argv[3]), (unsigned)std::atoi(argv[4]), std::atof(argv[5]), (unsigned)std::atoi(argv[6]),
                            std::atoi(argv[7]));
    check_device_overload<float>(14, 32, 1e-9, 2e-5);
    check_device_overload<double>(14, 64, 1e-22, 1e-10);
    check_device_overload<float>(2, 3, 1e-9, 2e-5);
    check_readme_name<float>(2e-4);
    check_readme_name<double>(1e-9);
    check_sizing_helpers();
    check_concurrent_workspaces();
    check_static_scratch(14, 900, 2e-5);
    std::printf(g_fail ? "api_surface: %d check(s) FAILED\n" : "api_surface: all checks passed\n", g_fail);
    return g_fail ? 1 : 0;
}
