// pcg_generic.cuh -- runtime-nx fallback kernel (any state size / horizon that fits one SM's shared memory).
//
// Used only when no specialised kernel (pcg_fused.cuh: nx <= 32, fp64 <= 25) is compiled for the requested nx.  One CTA per
// system, persistent over the batch; lambda, r, p, r~ live in shared memory, S and Phi^-1 are streamed
// from global memory (L2-resident after the first iteration) with row-consecutive, coalesced loads.
// Same recurrence and exit semantics as the fast path (reference include/pcg.cuh:118-212); same
// skip-the-padding-blocks column windows as reference include/utils.cuh:58-84.  No grid.sync.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pcg_rows.cuh"

namespace gbdpcg {

// shared memory: red [2][32] | xw [rows + 2*nx] (zero halo each side) | lam | res | p | rt [rows each]
template <typename T>
__host__ __device__ inline size_t generic_smem_bytes(uint32_t nx, uint32_t N)
{
    size_t rows = (size_t)nx * N;
    return sizeof(T) * (rows + 2 * (size_t)nx + 4 * rows + 64);
}

template <typename T, bool HAS_P>
__global__ void __launch_bounds__(1024, 1) pcg_generic_kernel(const KernelArgs<T> a, const uint32_t nx)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = (int)a.N, rows = N * (int)nx, nxx = (int)(nx * nx);
    T *red = reinterpret_cast<T *>(smem_raw);  // [2][32], first so that its 16-byte loads stay aligned
    T *xw = red + 64;                          // xw[nx + i] = x_i ; xw[0..nx) = xw[nx+rows ..) = 0
    T *lam = xw + rows + 2 * nx;
    T *res = lam + rows;
    T *pv = res + rows;
    T *rt = pv + rows;
    const int t = threadIdx.x, nthr = blockDim.x, lane = t & 31, warp = t >> 5;

    for (int i = t; i < (int)nx; i += nthr) { xw[i] = (T)0; xw[nx + rows + i] = (T)0; }
    for (int i = t; i < 64; i += nthr) red[i] = (T)0;
    __syncthreads();

    // y_i = sum_c M[b*3nx^2 + nx*c + r] * x[(b-1)*nx + c], c restricted to existing blocks
    auto row_mv = [&](const T *__restrict__ M, int i) -> T {
        const int b = i / (int)nx, r = i - b * (int)nx;
        const int c_lo = (b == 0) ? (int)nx : 0, c_hi = (b == N - 1) ? 2 * (int)nx : 3 * (int)nx;
        const T *m = M + (size_t)b * 3 * nxx + r;
        const T *w = xw + (size_t)b * nx;  // = x_{b-1} thanks to the nx-element left halo
        T acc0 = (T)0, acc1 = (T)0;
        int c = c_lo;
        for (; c + 1 < c_hi; c += 2) {
            acc0 += __ldg(m + (size_t)nx * c) * w[c];
            acc1 += __ldg(m + (size_t)nx * (c + 1)) * w[c + 1];
        }
        if (c < c_hi) acc0 += __ldg(m + (size_t)nx * c) * w[c];
        return acc0 + acc1;
    };
    auto all_sum = [&](T x, int slot) -> T {
        T ws = warp_allsum(x);
        T *rd = red + slot * 32;
        if (lane == 0) rd[warp] = ws;
        __syncthreads();
        return tree_sum_smem<T, 32>(rd);
    };

    for (uint32_t prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
        const T *S = a.S + (size_t)prob * a.mat_stride;
        const T *P = HAS_P ? a.Pinv + (size_t)prob * a.mat_stride : nullptr;
        const T *g = a.gamma + (size_t)prob * a.vec_stride;
        T *lg = a.lambda + (size_t)prob * a.vec_stride;

        for (int i = t; i < rows; i += nthr) { T l = lg[i]; lam[i] = l; xw[nx + i] = l; }
        __syncthreads();
        for (int i = t; i < rows; i += nthr) res[i] = __ldg(g + i) - row_mv(S, i);
        __syncthreads();
        T part = (T)0;
        if (HAS_P) {
            for (int i = t; i < rows; i += nthr) xw[nx + i] = res[i];
            __syncthreads();
            for (int i = t; i < rows; i += nthr) { T y = row_mv(P, i); rt[i] = y; pv[i] = y; part += res[i] * y; }
        } else {
            for (int i = t; i < rows; i += nthr) { T y = res[i]; rt[i] = y; pv[i] = y; part += y * y; }
        }
        T eta = all_sum(part, 1);

        uint32_t it = 0;
        bool hit_max = true;
        for (; it < a.max_iter; ++it) {
            for (int i = t; i < rows; i += nthr) xw[nx + i] = pv[i];
            __syncthreads();
            part = (T)0;
            for (int i = t; i < rows; i += nthr) { T u = row_mv(S, i); rt[i] = u; part += pv[i] * u; }  // rt holds upsilon for now
            const T v = all_sum(part, 0);
            const T alpha = eta / v;
            for (int i = t; i < rows; i += nthr) {
                lam[i] += alpha * pv[i];
                T rr = res[i] - alpha * rt[i];
                res[i] = rr;
                if (HAS_P) xw[nx + i] = rr;
            }
            part = (T)0;
            if (HAS_P) {
                __syncthreads();
                for (int i = t; i < rows; i += nthr) { T y = row_mv(P, i); rt[i] = y; part += res[i] * y; }
            } else {
                for (int i = t; i < rows; i += nthr) { T y = res[i]; rt[i] = y; part += y * y; }
            }
            const T eta_new = all_sum(part, 1);
            if (abs_t(eta_new) < a.tol) { ++it; hit_max = false; break; }
            const T beta = eta_new / eta;
            eta = eta_new;
            for (int i = t; i < rows; i += nthr) pv[i] = rt[i] + beta * pv[i];
            // pv[i], rt[i], res[i], lam[i] are only ever touched by thread (i mod nthr): no barrier needed here;
            // the xw rewrite at the top of the next iteration is ordered by the all_sum barrier above.
        }
        for (int i = t; i < rows; i += nthr) lg[i] = lam[i];
        if (t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        __syncthreads();
    }
}

}  // namespace gbdpcg
