// schur_form.cuh -- the steps either side of the solve: Schur-complement formation of
// S and gamma from stage-wise LQR data, and recovery of the primal step from lambda.
//
// Not in the reference tree: upstream (MPCGPU, cited at reference README.md:66-76) forms S, gamma on the GPU before it
// launches pcg<> (include/interface.cuh:92-103 is that seam); the reference's example input decodes to exactly this
// construction and gbd_pcg_b200/synth.py is its host restatement.  Sign convention of the example
// (S negative definite):  KKT  [G C^T; C 0][z; lambda] = [-g; -c],  G = blkdiag(Q_0, R_0, Q_1, R_1, ..., Q_{N-1}),
// C row-block 0 = [I],  row-block k+1 = [-A_k  -B_k  I]   =>   S = -C G^-1 C^T,  gamma = C G^-1 g - c:
//     S_00        = -Q_0^-1
//     S_{k+1,k+1} = -(A_k Q_k^-1 A_k^T + B_k R_k^-1 B_k^T + Q_{k+1}^-1)
//     S_{k+1,k}   =  A_k Q_k^-1            S_{k,k+1} = (A_k Q_k^-1)^T
//     gamma_0     =  Q_0^-1 q_0 - c_0
//     gamma_{k+1} = -A_k Q_k^-1 q_k - B_k R_k^-1 r_k + Q_{k+1}^-1 q_{k+1} - c_{k+1}
// and after the solve   x_k = -Q_k^-1 (q_k + lambda_k - A_k^T lambda_{k+1}),   u_k = -R_k^-1 (r_k - B_k^T lambda_{k+1}).
//
// Inputs are stage-major, every matrix column-major:  Q [batch][N][nx*nx], R [batch][N-1][nu*nu], A [batch][N-1][nx*nx],
// B [batch][N-1][nx*nu], q, c [batch][N][nx], r [batch][N-1][nu].  Output S in the solver's compressed layout
// (reference include/pcg.cuh:104-110), padding blocks zero.
//
// Two kernels, same arithmetic order (Gauss-Jordan without pivoting as glass::invertMatrix, GLASS/src/L3/inv.cuh:27-33;
// inner indices ascending):
//   schur_form_nx_kernel<T, NX>  compiled state sizes, nu <= nx: a CTA takes a chunk of block-rows of one system, stages
//     its stage data in shared memory (R_k, B_k padded to nx x nx with identity / zeros, which changes no value) and works
//     with the lane-per-columns mapping of pinv_build.cuh: all inversions, then A_k Q_k^-1 and B_k R_k^-1, then the
//     diagonal blocks and gamma; the S image leaves with coalesced 16-byte stores.
//   schur_form_kernel<T>         run-time nx, nu fallback: one CTA per (system, block-row).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pinv_build.cuh"   // gj_invert_cols, block_mv_bcast, copy_coalesced, pinv_cpl

namespace gbdpcg {

struct SchurArgs {
    const void *Q, *R, *A, *B, *q, *r, *c;
    void *S, *gamma, *Qinv, *Rinv;      // Qinv / Rinv optional by-products (for schur_recover)
    uint32_t nx, nu, N, batch;
#ifdef SCHUR_TRACE
    unsigned long long *trace;          // [8] accumulated clock64 per phase (thread 0 of every CTA), scripts/ubench/trace_schur.cu
#endif
};
#ifdef SCHUR_TRACE
#define SCHUR_MARK(slot) do { if (threadIdx.x == 0) { const long long now_ = clock64(); atomicAdd(a.trace + (slot), (unsigned long long)(now_ - tmark_)); tmark_ = now_; } } while (0)
#else
#define SCHUR_MARK(slot) do { } while (0)
#endif

// elements of shared memory the forming kernel needs
__host__ __device__ inline size_t schur_form_smem_elems(uint32_t nx, uint32_t nu)
{
    const size_t m = nx > nu ? nx : nu;
    return (size_t)5 * nx * nx + (size_t)nu * nu + (size_t)2 * nx * nu + 2 * m + 3 * nx + nu;
}

// in-place inverse of the n x n column-major matrix Am (shared memory); colv / rowv: n elements of scratch each
template <typename T>
__device__ __forceinline__ void gj_invert_smem(T *Am, int n, T *colv, T *rowv)
{
    const int t = threadIdx.x, nthr = blockDim.x;
    for (int p = 0; p < n; p++) {
        for (int e = t; e < n; e += nthr) { colv[e] = Am[p * n + e]; rowv[e] = Am[e * n + p]; }
        __syncthreads();
        const T pv_inv = (T)1 / colv[p];
        for (int e = t; e < n * n; e += nthr) {
            const int c = e / n, r = e - c * n;
            const T m = colv[r] * pv_inv;
            T v;
            if (r == p) v = (c == p) ? pv_inv : rowv[c] * pv_inv;
            else v = (c == p) ? -m : Am[e] - m * rowv[c];
            Am[e] = v;
        }
        __syncthreads();
    }
}

// Cm (m x n) = Am (m x k) * Bm (k x n) [or Bm^T when TRANS_B: Bm is n x k], column-major, inner index ascending
template <typename T, bool TRANS_B>
__device__ __forceinline__ void mm_smem(const T *Am, const T *Bm, T *Cm, int m, int k, int n)
{
    for (int e = threadIdx.x; e < m * n; e += blockDim.x) {
        const int c = e / m, r = e - c * m;
        T s = (T)0;
        for (int q = 0; q < k; q++) s += Am[q * m + r] * (TRANS_B ? Bm[q * n + c] : Bm[c * k + q]);
        Cm[e] = s;
    }
}

template <typename T>
__global__ void __launch_bounds__(128) schur_form_kernel(const SchurArgs a)
{
    extern __shared__ __align__(16) unsigned char schur_smem_raw[];
    const int nx = (int)a.nx, nu = (int)a.nu, N = (int)a.N;
    const int bb = nx * nx, uu = nu * nu, xu = nx * nu, mx = nx > nu ? nx : nu;
    T *Qk = reinterpret_cast<T *>(schur_smem_raw);   // Q_k^-1   (k = j-1)
    T *Qj = Qk + bb;                                  // Q_j^-1
    T *AQ = Qj + bb;                                  // A_k Q_k^-1          = S_{j,j-1}
    T *AQn = AQ + bb;                                 // A_j Q_j^-1          = S_{j+1,j}; its transpose is S_{j,j+1}
    T *Dg = AQn + bb;                                 // diagonal block accumulator
    T *Rk = Dg + bb;                                  // R_k^-1
    T *BR = Rk + uu;                                  // B_k R_k^-1          (nx x nu)
    T *Mt = BR + xu;                                  // staging for B_k     (A_k is staged in Dg)
    T *colv = Mt + xu, *rowv = colv + mx;
    T *v0 = rowv + mx, *v1 = v0 + nx, *v2 = v1 + nx, *vr = v2 + nx;   // q_k, q_j, c_j (nx each), r_k (nu)
    const int t = threadIdx.x, nthr = blockDim.x;
    const uint32_t items = a.batch * a.N;

    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const uint32_t prob = item / a.N;
        const int j = (int)(item - prob * a.N), k = j - 1;
        const T *Q = static_cast<const T *>(a.Q) + ((size_t)prob * N) * bb;
        const T *R = static_cast<const T *>(a.R) + ((size_t)prob * (N - 1)) * uu;
        const T *A = static_cast<const T *>(a.A) + ((size_t)prob * (N - 1)) * bb;
        const T *B = static_cast<const T *>(a.B) + ((size_t)prob * (N - 1)) * xu;
        const T *q = static_cast<const T *>(a.q) + ((size_t)prob * N) * nx;
        const T *r = static_cast<const T *>(a.r) + ((size_t)prob * (N - 1)) * nu;
        const T *c = static_cast<const T *>(a.c) + ((size_t)prob * N) * nx;
        T *Sr = static_cast<T *>(a.S) + ((size_t)prob * N + j) * 3 * bb;
        T *gam = static_cast<T *>(a.gamma) + ((size_t)prob * N + j) * nx;

        // ---- loads ----
        for (int e = t; e < bb; e += nthr) { Qj[e] = Q[(size_t)j * bb + e]; if (k >= 0) Qk[e] = Q[(size_t)k * bb + e]; }
        if (k >= 0) for (int e = t; e < uu; e += nthr) Rk[e] = R[(size_t)k * uu + e];
        for (int e = t; e < nx; e += nthr) { v1[e] = q[(size_t)j * nx + e]; v2[e] = c[(size_t)j * nx + e]; if (k >= 0) v0[e] = q[(size_t)k * nx + e]; }
        if (k >= 0) for (int e = t; e < nu; e += nthr) vr[e] = r[(size_t)k * nu + e];
        __syncthreads();

        // ---- inverses ----
        gj_invert_smem<T>(Qj, nx, colv, rowv);
        if (k >= 0) { gj_invert_smem<T>(Qk, nx, colv, rowv); gj_invert_smem<T>(Rk, nu, colv, rowv); }
        if (a.Qinv) for (int e = t; e < bb; e += nthr) static_cast<T *>(a.Qinv)[((size_t)prob * N + j) * bb + e] = Qj[e];
        if (a.Rinv && k >= 0) for (int e = t; e < uu; e += nthr) static_cast<T *>(a.Rinv)[((size_t)prob * (N - 1) + k) * uu + e] = Rk[e];

        // ---- right block: S_{j,j+1} = (A_j Q_j^-1)^T ----
        if (j + 1 < N) {
            // AQn = A_j * Qj, reading A_j straight from global memory
            const T *Aj = A + (size_t)j * bb;
            for (int e = t; e < bb; e += nthr) {
                const int cc = e / nx, rr = e - cc * nx;
                T s = (T)0;
                for (int p = 0; p < nx; p++) s += Aj[p * nx + rr] * Qj[cc * nx + p];
                AQn[e] = s;
            }
            __syncthreads();
            for (int e = t; e < bb; e += nthr) { const int cc = e / nx, rr = e - cc * nx; Sr[2 * bb + e] = AQn[rr * nx + cc]; }
        } else {
            for (int e = t; e < bb; e += nthr) Sr[2 * bb + e] = (T)0;
        }

        if (k < 0) {
            // ---- block-row 0:  S_00 = -Q_0^-1, gamma_0 = Q_0^-1 q_0 - c_0 ----
            for (int e = t; e < bb; e += nthr) { Sr[e] = (T)0; Sr[bb + e] = -Qj[e]; }
            for (int e = t; e < nx; e += nthr) {
                T s = (T)0;
                for (int p = 0; p < nx; p++) s += Qj[p * nx + e] * v1[p];
                gam[e] = s - v2[e];
            }
        } else {
            // ---- block-row j = k+1 ----
            const T *Ak = A + (size_t)k * bb, *Bk = B + (size_t)k * xu;
            for (int e = t; e < bb; e += nthr) Dg[e] = Ak[e];                  // stage A_k
            for (int e = t; e < xu; e += nthr) Mt[e] = Bk[e];                  // stage B_k
            __syncthreads();
            mm_smem<T, false>(Dg, Qk, AQ, nx, nx, nx);                         // AQ = A_k Q_k^-1
            mm_smem<T, false>(Mt, Rk, BR, nx, nu, nu);                         // BR = B_k R_k^-1
            __syncthreads();
            for (int e = t; e < bb; e += nthr) Sr[e] = AQ[e];                  // left block
            // diag = -((AQ A_k^T + BR B_k^T) + Q_j^-1);  A_k is still in Dg, B_k in Mt
            for (int e = t; e < bb; e += nthr) {
                const int cc = e / nx, rr = e - cc * nx;
                T s = (T)0;
                for (int p = 0; p < nx; p++) s += AQ[p * nx + rr] * Dg[p * nx + cc];
                T s2 = (T)0;
                for (int p = 0; p < nu; p++) s2 += BR[p * nx + rr] * Mt[p * nx + cc];
                Sr[bb + e] = -((s + s2) + Qj[e]);
            }
            // gamma_j = ((Q_j^-1 q_j - AQ q_k) - BR r_k) - c_j
            for (int e = t; e < nx; e += nthr) {
                T s = (T)0, s1 = (T)0, s2 = (T)0;
                for (int p = 0; p < nx; p++) { s += Qj[p * nx + e] * v1[p]; s1 += AQ[p * nx + e] * v0[p]; }
                for (int p = 0; p < nu; p++) s2 += BR[p * nx + e] * vr[p];
                gam[e] = ((s - s1) - s2) - v2[e];
            }
        }
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// specialised kernel
// ---------------------------------------------------------------------------------------------------------------------
// shared memory (elements): Qs, Rs, As, Bs [chunk+1][NX*NX] | img [chunk][3 NX*NX] | scratch | q, r [chunk+1][NX] | c [chunk][NX] | r compact [chunk+1][NX]
template <typename T>
__host__ __device__ inline size_t schur_nx_smem_bytes(int nx, int chunk, int threads)
{
    const int groups = (threads / 32) * (32 / (nx / pinv_cpl(nx)));
    const size_t w = (size_t)chunk + 1, bb = (size_t)nx * nx;
    return (4 * w * bb + 3 * (size_t)chunk * bb + 3 * w * nx + (size_t)chunk * nx + (size_t)groups * 16) * sizeof(T) + 32;
}

template <typename T, int NX>
__global__ void __launch_bounds__(256, (sizeof(T) == 4 ? 2 : 1)) schur_form_nx_kernel(const SchurArgs a, const uint32_t chunk_, const uint32_t nchunks)
{
    constexpr int BB = NX * NX, BSZ = 3 * BB;
    constexpr int CPL = pinv_cpl(NX), LPB = NX / CPL, GPW = 32 / LPB;
    extern __shared__ __align__(16) unsigned char schur_smem_raw[];
    const int chunk = (int)chunk_, W = chunk + 1, N = (int)a.N, nu = (int)a.nu;
    T *Qs = reinterpret_cast<T *>(schur_smem_raw);
    T *Rs = Qs + (size_t)W * BB, *As = Rs + (size_t)W * BB, *Bs = As + (size_t)W * BB;
    T *img = Bs + (size_t)W * BB;
    const int t = threadIdx.x, nthr = blockDim.x, lane = t & 31, warp = t >> 5;
    const int gl = lane / LPB, li = lane - gl * LPB;
    const bool lane_on = gl < GPW;
    const int group = warp * GPW + gl, ngroups = (nthr >> 5) * GPW;
    T *scr_all = img + (size_t)chunk * BSZ;                     // 16-byte aligned: read with vector loads
    T *vq = scr_all + (size_t)ngroups * 16, *vr = vq + (size_t)W * NX, *vc = vr + (size_t)W * NX, *vrc = vc + (size_t)chunk * NX;
    T *scr = scr_all + (lane_on ? group : 0) * 16;
    const int c0 = li * CPL;                                    // first column this lane owns
    // phase-3 mapping: C3 = 1 column per lane, NX lanes per block-row
    constexpr int C3 = 1, LPB3 = NX / C3, GPW3 = 32 / LPB3;
    const int gl3 = lane / LPB3, c3 = (lane - gl3 * LPB3) * C3;
    const bool lane_on3 = gl3 < GPW3;
    const int group3 = warp * GPW3 + gl3, ngroups3 = (nthr >> 5) * GPW3;
    const uint32_t items = a.batch * nchunks;
    bool bulk_pending = false;

    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const uint32_t prob = item / nchunks;
        const int k0 = (int)(item - prob * nchunks) * chunk;
        const int cnt = min(chunk, N - k0);
        const T *Q = static_cast<const T *>(a.Q) + ((size_t)prob * N) * BB;
        const T *R = static_cast<const T *>(a.R) + ((size_t)prob * (N - 1)) * nu * nu;
        const T *A = static_cast<const T *>(a.A) + ((size_t)prob * (N - 1)) * BB;
        const T *B = static_cast<const T *>(a.B) + ((size_t)prob * (N - 1)) * NX * nu;
        const T *q = static_cast<const T *>(a.q) + ((size_t)prob * N) * NX;
        const T *r = static_cast<const T *>(a.r) + ((size_t)prob * (N - 1)) * nu;
        const T *c = static_cast<const T *>(a.c) + ((size_t)prob * N) * NX;
        // slot i <-> knot kk = k0 - 1 + i;  row lb <-> block-row j = k0 + lb uses slot lb (knot j-1) and slot lb+1 (knot j)
        const int i_lo = (k0 == 0) ? 1 : 0;                    // first slot that exists
        auto knot = [&](int i) { return k0 - 1 + i; };

#ifdef SCHUR_TRACE
        long long tmark_ = clock64();
#endif
        if (bulk_pending) {                                    // the previous item's bulk stores must have finished reading Qs / img
            if (t == 0) pv_bulk_wait_read();
            __syncthreads();
            bulk_pending = false;
        }
        // ---- phase 0: stage the chunk's data ----
        // (every piece is an asynchronous cp.async copy: all of them are in flight together, one wait)
        async_copy<T>(Qs + (size_t)i_lo * BB, Q + (size_t)knot(i_lo) * BB, (cnt + 1 - i_lo) * BB);
        {
            const int hi = min(cnt, N - 2 - knot(0));           // last slot whose A / B / R exists
            const int ns = hi + 1 - i_lo;                       // slots with A / B / R
            T *tmpR = img, *tmpB = img + (size_t)W * BB;        // compact copies parked in the (still unused) output image
            if (ns > 0) {
                async_copy<T>(As + (size_t)i_lo * BB, A + (size_t)knot(i_lo) * BB, ns * BB);
                async_copy<T>(tmpR, R + (size_t)knot(i_lo) * nu * nu, ns * nu * nu);
                async_copy<T>(tmpB, B + (size_t)knot(i_lo) * NX * nu, ns * NX * nu);
                async_copy<T>(vrc, r + (size_t)knot(i_lo) * nu, ns * nu);
            }
            async_copy<T>(vq + i_lo * NX, q + (size_t)knot(i_lo) * NX, (cnt + 1 - i_lo) * NX);
            async_copy<T>(vc, c + (size_t)k0 * NX, cnt * NX);
            async_copy_wait();
            __syncthreads();
            for (int e = t; e < ns * NX; e += nthr) {
                const int s_ = e / NX, w = e - s_ * NX;
                vr[(i_lo + s_) * NX + w] = (w < nu) ? vrc[s_ * nu + w] : (T)0;
            }
            // expand R_k, B_k to nx x nx (identity / zero padding), shared -> shared
            for (int e = t; e < ns * BB; e += nthr) {
                const int s_ = e / BB, w = e - s_ * BB, cc = w / NX, rr = w - cc * NX;
                Rs[(size_t)(i_lo + s_) * BB + w] = (rr < nu && cc < nu) ? tmpR[s_ * nu * nu + cc * nu + rr] : ((rr == cc) ? (T)1 : (T)0);
                Bs[(size_t)(i_lo + s_) * BB + w] = (cc < nu) ? tmpB[s_ * NX * nu + cc * NX + rr] : (T)0;
            }
        }
        __syncthreads();
        SCHUR_MARK(0);

        // ---- phase 1: Q_kk^-1 (slots i_lo..cnt) and padded R_kk^-1 (slots i_lo..cnt-1, kk <= N-2), in place ----
        const int hi_ab = min(cnt, N - 1 - k0);                 // last slot whose A exists (knot <= N-2)
        const int nQ = cnt + 1 - i_lo, nR = max(0, min(cnt - 1, hi_ab) - i_lo + 1);
        for (int base = 0; base < nQ + nR; base += ngroups) {
            const int ti = base + group;
            const bool isR = ti >= nQ;
            const int i = i_lo + (isR ? ti - nQ : ti);
            const bool valid = lane_on && ti < nQ + nR;
            if (!__any_sync(0xffffffffu, valid)) continue;      // warp-uniform
            T *blk = (isR ? Rs : Qs) + (size_t)(valid ? i : i_lo) * BB + c0 * NX;
            T col[CPL][NX];
#pragma unroll
            for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                for (int rr = 0; rr < NX; rr++) col[cc][rr] = valid ? blk[cc * NX + rr] : ((rr == c0 + cc) ? (T)1 : (T)0);
            // R blocks are padded with an identity tail: nu pivots suffice; the warp runs as many as its longest block needs
            const int nsteps = __reduce_max_sync(0xffffffffu, valid ? (isR ? nu : NX) : 0);
            gj_invert_cols<T, NX, CPL>(col, valid, li, scr, nsteps);
            if (valid) {
#pragma unroll
                for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                    for (int rr = 0; rr < NX; rr++) blk[cc * NX + rr] = col[cc][rr];
            }
        }
        __syncthreads();
        SCHUR_MARK(1);

        // ---- phase 2: AQ_kk = A_kk Q_kk^-1 (left block of row kk+1, transposed: right block of row kk);
        //               BR_kk = B_kk R_kk^-1 (parked in the diagonal slot of row kk+1) ----
        const int nA = max(0, hi_ab - i_lo + 1), nB = nR;
        for (int base = 0; base < nA + nB; base += ngroups) {
            const int ti = base + group;
            const bool isB = ti >= nA;
            const int i = i_lo + (isB ? ti - nA : ti);
            const bool valid = lane_on && ti < nA + nB;
            if (valid) {
                T x[CPL][NX], w[CPL][NX];
                const T *X = (isB ? Rs : Qs) + (size_t)i * BB + c0 * NX;
#pragma unroll
                for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                    for (int rr = 0; rr < NX; rr++) x[cc][rr] = X[cc * NX + rr];
                block_mv_bcast<T, NX, CPL>((isB ? Bs : As) + (size_t)i * BB, x, w);
                if (isB) {
#pragma unroll
                    for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                        for (int rr = 0; rr < NX; rr++) img[(size_t)i * BSZ + BB + (c0 + cc) * NX + rr] = w[cc][rr];
                } else {
                    if (i <= cnt - 1) {
#pragma unroll
                        for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                            for (int rr = 0; rr < NX; rr++) img[(size_t)i * BSZ + (c0 + cc) * NX + rr] = w[cc][rr];
                    }
                    if (i >= 1) {
#pragma unroll
                        for (int cc = 0; cc < CPL; cc++)
#pragma unroll
                            for (int rr = 0; rr < NX; rr++) img[(size_t)(i - 1) * BSZ + 2 * BB + rr * NX + (c0 + cc)] = w[cc][rr];
                    }
                }
            }
        }
        __syncthreads();
        SCHUR_MARK(2);

        // ---- phase 3: diagonal blocks and gamma, row by row (one column per lane: a row keeps NX lanes busy, so the
        //      chunk's rows fill the CTA in one round) ----
        for (int base = 0; base < cnt; base += ngroups3) {
            const int lb = base + group3;
            const bool valid = lane_on3 && lb < cnt;
            if (!__any_sync(0xffffffffu, valid)) continue;      // warp-uniform
            const int lbv = valid ? lb : 0, j = k0 + lbv;
            const T *Qj = Qs + (size_t)(lbv + 1) * BB;
            T *row = img + (size_t)lbv * BSZ;
            T out[C3][NX];
#pragma unroll
            for (int cc = 0; cc < C3; cc++)
#pragma unroll
                for (int rr = 0; rr < NX; rr++) out[cc][rr] = -Qj[(c3 + cc) * NX + rr];
            T gam[C3];
#pragma unroll
            for (int cc = 0; cc < C3; cc++) {
                T s = (T)0;
#pragma unroll
                for (int p = 0; p < NX; p++) s += Qj[p * NX + c3 + cc] * vq[(lbv + 1) * NX + p];
                gam[cc] = s;
            }
            if (valid && j > 0) {
                T v[C3][NX], w2[C3][NX];
#pragma unroll
                for (int cc = 0; cc < C3; cc++)
#pragma unroll
                    for (int p = 0; p < NX; p++) v[cc][p] = As[(size_t)lbv * BB + p * NX + c3 + cc];       // row c0+cc of A_k
                block_mv_bcast<T, NX, C3>(row, v, out);                                                  // AQ A_k^T
#pragma unroll
                for (int cc = 0; cc < C3; cc++)
#pragma unroll
                    for (int p = 0; p < NX; p++) v[cc][p] = Bs[(size_t)lbv * BB + p * NX + c3 + cc];       // row c0+cc of B_k (padded)
                block_mv_bcast<T, NX, C3>(row + BB, v, w2);                                              // BR B_k^T
#pragma unroll
                for (int cc = 0; cc < C3; cc++)
#pragma unroll
                    for (int rr = 0; rr < NX; rr++) out[cc][rr] = -((out[cc][rr] + w2[cc][rr]) + Qj[(c3 + cc) * NX + rr]);
#pragma unroll
                for (int cc = 0; cc < C3; cc++) {
                    T s1 = (T)0, s2 = (T)0;
#pragma unroll
                    for (int p = 0; p < NX; p++) {
                        s1 += row[p * NX + c3 + cc] * vq[lbv * NX + p];
                        s2 += row[BB + p * NX + c3 + cc] * vr[lbv * NX + p];
                    }
                    gam[cc] = (gam[cc] - s1) - s2;
                }
            }
            __syncwarp();          // BR (parked in the diagonal slot) has been read by the whole group
            if (valid) {
#pragma unroll
                for (int cc = 0; cc < C3; cc++) {
#pragma unroll
                    for (int rr = 0; rr < NX; rr++) {
                        row[BB + (c3 + cc) * NX + rr] = out[cc][rr];
                        if (j == 0) row[(c3 + cc) * NX + rr] = (T)0;
                        if (j == N - 1) row[2 * BB + (c3 + cc) * NX + rr] = (T)0;
                    }
                    static_cast<T *>(a.gamma)[((size_t)prob * N + j) * NX + c3 + cc] = gam[cc] - vc[lbv * NX + c3 + cc];
                }
            }
        }
        __syncthreads();
        SCHUR_MARK(3);

        // ---- phase 4: results (S image and Q^-1 leave through the TMA engine when 16-byte aligned) ----
        {
            T *Sdst = static_cast<T *>(a.S) + ((size_t)prob * N + k0) * BSZ;
            T *Qdst = a.Qinv ? static_cast<T *>(a.Qinv) + ((size_t)prob * N + k0) * BB : nullptr;
            const bool bulk = (sizeof(T) * BB) % 16 == 0 && (reinterpret_cast<uintptr_t>(Sdst) & 15u) == 0 &&
                              (reinterpret_cast<uintptr_t>(Qdst) & 15u) == 0;
            if (bulk) {
                pv_fence_async();          // this thread's generic-proxy writes of the image -> visible to the async proxy
                __syncthreads();
                if (t == 0) {
                    pv_bulk_s2g(Sdst, pv_smem_u32(img), (uint32_t)(cnt * BSZ * (int)sizeof(T)));
                    if (Qdst) pv_bulk_s2g(Qdst, pv_smem_u32(Qs + BB), (uint32_t)(cnt * BB * (int)sizeof(T)));
                }
                bulk_pending = true;
            } else {
                copy_coalesced<T>(Sdst, img, cnt * BSZ, false);
                if (Qdst) copy_coalesced<T>(Qdst, Qs + BB, cnt * BB, false);
            }
        }
        if (a.Rinv) {
            T *Ro = static_cast<T *>(a.Rinv) + ((size_t)prob * (N - 1)) * nu * nu;
            for (int e = t; e < cnt * nu * nu; e += nthr) {
                const int i = e / (nu * nu), w = e - i * nu * nu, cc = w / nu, rr = w - cc * nu, kk = knot(i);
                if (i >= i_lo && kk <= N - 2) Ro[(size_t)kk * nu * nu + w] = Rs[(size_t)i * BB + cc * NX + rr];
            }
        }
        __syncthreads();
        SCHUR_MARK(4);
    }
    if (bulk_pending && t == 0) pv_bulk_wait_read();           // shared memory must outlive the last bulk store
}

// x_k = -Q_k^-1 (q_k + lambda_k - A_k^T lambda_{k+1}),  u_k = -R_k^-1 (r_k - B_k^T lambda_{k+1});  one CTA per (system, k)
template <typename T>
__global__ void __launch_bounds__(64) schur_recover_kernel(const T *__restrict__ Qinv, const T *__restrict__ Rinv, const T *__restrict__ A,
                                                           const T *__restrict__ B, const T *__restrict__ q, const T *__restrict__ r,
                                                           const T *__restrict__ lambda, T *__restrict__ x, T *__restrict__ u,
                                                           uint32_t nx_, uint32_t nu_, uint32_t N_, uint32_t batch)
{
    extern __shared__ __align__(16) unsigned char schur_smem_raw[];
    const int nx = (int)nx_, nu = (int)nu_, N = (int)N_;
    T *tx = reinterpret_cast<T *>(schur_smem_raw), *tu = tx + nx;
    const int t = threadIdx.x, nthr = blockDim.x;
    const uint32_t items = batch * N_;
    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const uint32_t prob = item / N_;
        const int k = (int)(item - prob * N_);
        const T *lam = lambda + (size_t)prob * N * nx;
        const bool has_next = k + 1 < N;
        const T *Ak = A + ((size_t)prob * (N - 1) + k) * nx * nx, *Bk = B + ((size_t)prob * (N - 1) + k) * nx * nu;
        for (int e = t; e < nx; e += nthr) {
            T s = (T)0;
            if (has_next) for (int p = 0; p < nx; p++) s += Ak[e * nx + p] * lam[(size_t)(k + 1) * nx + p];      // (A_k^T lambda_{k+1})[e]
            tx[e] = (q[((size_t)prob * N + k) * nx + e] + lam[(size_t)k * nx + e]) - s;
        }
        if (has_next) for (int e = t; e < nu; e += nthr) {
            T s = (T)0;
            for (int p = 0; p < nx; p++) s += Bk[e * nx + p] * lam[(size_t)(k + 1) * nx + p];                     // (B_k^T lambda_{k+1})[e]
            tu[e] = r[((size_t)prob * (N - 1) + k) * nu + e] - s;
        }
        __syncthreads();
        const T *Qi = Qinv + ((size_t)prob * N + k) * nx * nx;
        for (int e = t; e < nx; e += nthr) {
            T s = (T)0;
            for (int p = 0; p < nx; p++) s += Qi[p * nx + e] * tx[p];
            x[((size_t)prob * N + k) * nx + e] = -s;
        }
        if (has_next) {
            const T *Ri = Rinv + ((size_t)prob * (N - 1) + k) * nu * nu;
            for (int e = t; e < nu; e += nthr) {
                T s = (T)0;
                for (int p = 0; p < nu; p++) s += Ri[p * nu + e] * tu[p];
                u[((size_t)prob * (N - 1) + k) * nu + e] = -s;
            }
        }
        __syncthreads();
    }
}

// Warp-per-(system, knot) recovery for nx, nu <= 32: lane e owns row e; the intermediate vectors are exchanged with
// shuffles, no shared memory, no block barrier.  Same arithmetic order as schur_recover_kernel.
template <typename T>
__global__ void __launch_bounds__(256) schur_recover_warp_kernel(const T *__restrict__ Qinv, const T *__restrict__ Rinv,
                                                                 const T *__restrict__ A, const T *__restrict__ B, const T *__restrict__ q,
                                                                 const T *__restrict__ r, const T *__restrict__ lambda, T *__restrict__ x,
                                                                 T *__restrict__ u, uint32_t nx_, uint32_t nu_, uint32_t N_, uint32_t batch)
{
    const int nx = (int)nx_, nu = (int)nu_, N = (int)N_;
    const int lane = threadIdx.x & 31;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t items = batch * N_;
    for (uint32_t item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; item < items; item += warps) {
        const uint32_t prob = item / N_;
        const int k = (int)(item - prob * N_);
        const bool has_next = k + 1 < N;
        const T *lam = lambda + (size_t)prob * N * nx;
        const T ln = (has_next && lane < nx) ? lam[(size_t)(k + 1) * nx + lane] : (T)0;      // lambda_{k+1}[lane]
        T tx = (T)0, tu = (T)0;
        if (has_next) {
            const T *Ak = A + ((size_t)prob * (N - 1) + k) * nx * nx, *Bk = B + ((size_t)prob * (N - 1) + k) * nx * nu;
            T s = (T)0, s2 = (T)0;
            for (int p = 0; p < nx; p++) {
                const T lp = __shfl_sync(0xffffffffu, ln, p);
                if (lane < nx) s += Ak[lane * nx + p] * lp;                                   // (A_k^T lambda_{k+1})[lane]
                if (lane < nu) s2 += Bk[lane * nx + p] * lp;                                  // (B_k^T lambda_{k+1})[lane]
            }
            tx = s; tu = s2;
        }
        if (lane < nx) tx = (q[((size_t)prob * N + k) * nx + lane] + lam[(size_t)k * nx + lane]) - tx;
        if (has_next && lane < nu) tu = r[((size_t)prob * (N - 1) + k) * nu + lane] - tu;
        const T *Qi = Qinv + ((size_t)prob * N + k) * nx * nx;
        T sx = (T)0;
        for (int p = 0; p < nx; p++) {
            const T tp = __shfl_sync(0xffffffffu, tx, p);
            if (lane < nx) sx += Qi[p * nx + lane] * tp;
        }
        if (lane < nx) x[((size_t)prob * N + k) * nx + lane] = -sx;
        if (has_next) {
            const T *Ri = Rinv + ((size_t)prob * (N - 1) + k) * nu * nu;
            T su = (T)0;
            for (int p = 0; p < nu; p++) {
                const T tp = __shfl_sync(0xffffffffu, tu, p);
                if (lane < nu) su += Ri[p * nu + lane] * tp;
            }
            if (lane < nu) u[((size_t)prob * (N - 1) + k) * nu + lane] = -su;
        }
    }
}

// Staged form of the warp-per-(system, knot) recovery: the knot's four matrices reach the warp's own shared-memory slice by cp.async
// (16-byte pieces where the source allows; every byte of Q^-1, R^-1, A, B is requested once, fully coalesced) and the NEXT TWO items of the
// warp are in flight while the current one is computed (three slices per warp).  A_k^T lambda reads the slice with a lane stride of nx
// words, conflict-free for the even state sizes; the global-memory form above pays one L1 wavefront per 128-byte line of that stride.
// Arithmetic and its order are those of schur_recover_warp_kernel (bit-identical results).
template <typename T>
__host__ __device__ inline uint32_t schur_recover_slice_elems(uint32_t nx, uint32_t nu)
{
    const uint32_t al = 16 / (uint32_t)sizeof(T);
    auto up = [al](uint32_t n) { return (n + al - 1) / al * al; };
    return 2 * up(nx * nx) + up(nx * nu) + up(nu * nu);
}

// COUNT > 0: the element count at compile time (loops unrolled).  PIECE = bytes per cp.async: the largest of 16 / 8 / sizeof(T) that
// divides both the source address and COUNT * sizeof(T) when COUNT is known, else chosen from the address with an element-sized tail.
template <typename T, int COUNT>
__device__ __forceinline__ void warp_async_copy(T *dst, const T *src, int count, int lane)
{
    const uint32_t d0 = pv_smem_u32(dst);                        // 16-byte aligned by construction
    const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
    const char *s0 = reinterpret_cast<const char *>(src);
    if constexpr (COUNT > 0) {
        constexpr int BYTES = COUNT * (int)sizeof(T);
        if (BYTES % 16 == 0 && (sa & 15u) == 0) {
#pragma unroll
            for (int q0 = 0; q0 < BYTES / 16; q0 += 32) {
                const int q = q0 + lane;
                if (q < BYTES / 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * (uint32_t)q), "l"(s0 + 16 * q) : "memory");
            }
        } else if (BYTES % 8 == 0 && (sa & 7u) == 0) {
#pragma unroll
            for (int q0 = 0; q0 < BYTES / 8; q0 += 32) {
                const int q = q0 + lane;
                if (q < BYTES / 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + 8u * (uint32_t)q), "l"(s0 + 8 * q) : "memory");
            }
        } else {
#pragma unroll
            for (int q0 = 0; q0 < COUNT; q0 += 32) {
                const int q = q0 + lane;
                if (q < COUNT) {
                    if constexpr (sizeof(T) == 4)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4u * (uint32_t)q), "l"(src + q) : "memory");
                    else
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + 8u * (uint32_t)q), "l"(src + q) : "memory");
                }
            }
        }
    } else {
        const int bytes = count * (int)sizeof(T);
        int done = 0;
        if ((sa & 15u) == 0) {
            const int n16 = bytes / 16;
            for (int q = lane; q < n16; q += 32)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * (uint32_t)q), "l"(s0 + 16 * (size_t)q) : "memory");
            done = n16 * 16;
        } else if ((sa & 7u) == 0) {
            const int n8 = bytes / 8;
            for (int q = lane; q < n8; q += 32)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + 8u * (uint32_t)q), "l"(s0 + 8 * (size_t)q) : "memory");
            done = n8 * 8;
        }
        for (int q = done / (int)sizeof(T) + lane; q < count; q += 32) {
            if constexpr (sizeof(T) == 4)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4u * (uint32_t)q), "l"(src + q) : "memory");
            else
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + 8u * (uint32_t)q), "l"(src + q) : "memory");
        }
    }
}

// NX, NU > 0: state / input size at compile time (inner loops and the copy loops unrolled, immediate offsets -- the run-time form,
// NX = NU = 0, is bound by instruction issue: 559 warp instructions per knot against ~250 here).
template <typename T, int NX, int NU>
__global__ void __launch_bounds__(256, (sizeof(T) == 4 ? 4 : 2)) schur_recover_staged_kernel(const T *__restrict__ Qinv, const T *__restrict__ Rinv,
                                                                   const T *__restrict__ A, const T *__restrict__ B, const T *__restrict__ q,
                                                                   const T *__restrict__ r, const T *__restrict__ lambda, T *__restrict__ x,
                                                                   T *__restrict__ u, uint32_t nx_, uint32_t nu_, uint32_t N_, uint32_t batch)
{
    extern __shared__ __align__(16) unsigned char schur_smem_raw[];
    const int nx = NX > 0 ? NX : (int)nx_, nu = NU > 0 ? NU : (int)nu_, N = (int)N_;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t al = 16 / (uint32_t)sizeof(T);
    const uint32_t e_bb = ((uint32_t)(nx * nx) + al - 1) / al * al, e_xu = ((uint32_t)(nx * nu) + al - 1) / al * al;
    const uint32_t slice = schur_recover_slice_elems<T>((uint32_t)nx, (uint32_t)nu);
    T *mine = reinterpret_cast<T *>(schur_smem_raw) + (size_t)wib * 3 * slice;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t items = batch * N_;
    const int lx = lane < nx ? lane : 0, lu = lane < nu ? lane : 0;
    // the item's matrices -> slice `buf` (asynchronous); its vector elements -> registers (used one loop trip later)
    auto issue = [&](uint32_t item, int buf, T &ln, T &qk, T &lk, T &rk) {
        const uint32_t prob = item / N_;
        const int k = (int)(item - prob * N_);
        T *Qs = mine + (size_t)buf * slice, *As = Qs + e_bb, *Bs = As + e_bb, *Rs = Bs + e_xu;
        warp_async_copy<T, NX * NX>(Qs, Qinv + (size_t)item * (nx * nx), nx * nx, lane);
        const T *lam = lambda + (size_t)item * nx;
        const bool has_next = k + 1 < N;
        if (has_next) {
            const size_t kk = (size_t)item - prob;                                            // prob * (N - 1) + k
            warp_async_copy<T, NX * NX>(As, A + kk * (nx * nx), nx * nx, lane);
            warp_async_copy<T, NX * NU>(Bs, B + kk * (nx * nu), nx * nu, lane);
            warp_async_copy<T, NU * NU>(Rs, Rinv + kk * (nu * nu), nu * nu, lane);
            rk = r[kk * nu + lu];
            ln = lam[nx + lx];                                                                // lambda_{k+1}[lane]
        } else {
            rk = (T)0;
            ln = (T)0;
        }
        qk = q[(size_t)item * nx + lx];                   // (q_k and lambda_k are added when they are USED: an add here would wait for both loads)
        lk = lam[lx];
    };
    uint32_t item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int buf = 0;
    // three slices per warp: the item being computed and the next TWO in flight (one ahead left the warp waiting on HBM for
    // about as long as it computes: 150 -> see DESIGN 3.6)
    T ln = (T)0, qk = (T)0, lk = (T)0, rk = (T)0, ln_1 = (T)0, qk_1 = (T)0, lk_1 = (T)0, rk_1 = (T)0;
    if (item < items) issue(item, 0, ln, qk, lk, rk);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (item + warps < items) issue(item + warps, 1, ln_1, qk_1, lk_1, rk_1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (; item < items; item += warps) {
        T ln_n = (T)0, qk_n = (T)0, lk_n = (T)0, rk_n = (T)0;
        const int buf2 = buf >= 1 ? buf - 1 : 2;                  // (buf + 2) mod 3
        if ((unsigned long long)item + 2ull * warps < items) issue(item + 2 * warps, buf2, ln_n, qk_n, lk_n, rk_n);
        asm volatile("cp.async.commit_group;" ::: "memory");
        const uint32_t prob = item / N_;
        const int k = (int)(item - prob * N_);
        const bool has_next = k + 1 < N;
        asm volatile("cp.async.wait_group 2;" ::: "memory");      // this item's copies (the group two before the one just committed)
        __syncwarp();
        const T *Qs = mine + (size_t)buf * slice, *As = Qs + e_bb, *Bs = As + e_bb, *Rs = Bs + e_xu;
        T tx = (T)0, tu = (T)0;
        if (has_next) {
            T s = (T)0, s2 = (T)0;
#pragma unroll
            for (int p = 0; p < nx; p++) {
                const T lp = __shfl_sync(0xffffffffu, ln, p);
                s += As[lx * nx + p] * lp;                                                    // (A_k^T lambda_{k+1})[lane]
                s2 += Bs[lu * nx + p] * lp;                                                   // (B_k^T lambda_{k+1})[lane]
            }
            tx = s; tu = s2;
        }
        tx = (qk + lk) - tx;
        tu = rk - tu;
        T sx = (T)0;
#pragma unroll
        for (int p = 0; p < nx; p++) sx += Qs[p * nx + lx] * __shfl_sync(0xffffffffu, tx, p);
        if (lane < nx) x[(size_t)item * nx + lane] = -sx;
        if (has_next) {
            T su = (T)0;
#pragma unroll
            for (int p = 0; p < nu; p++) su += Rs[p * nu + lu] * __shfl_sync(0xffffffffu, tu, p);
            if (lane < nu) u[((size_t)item - prob) * nu + lane] = -su;
        }
        __syncwarp();                                             // the slice is free again
        ln = ln_1; qk = qk_1; lk = lk_1; rk = rk_1;
        ln_1 = ln_n; qk_1 = qk_n; lk_1 = lk_n; rk_1 = rk_n;
        buf = buf == 2 ? 0 : buf + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

}  // namespace gbdpcg
