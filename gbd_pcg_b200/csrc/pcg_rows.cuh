// pcg_rows.cuh -- the register-row PCG kernel for sm_100a.
//
// One thread owns one row of the (N*nx)-dimensional system; one CTA owns a contiguous run of
// block-rows (knot points); one thread-block cluster (1..16 CTAs) owns one whole system and stays
// resident for the complete solve, looping over a batch of systems.  Compared with the reference
// kernel (include/pcg.cuh:54-218: one 64-thread CTA per block-row, vectors exchanged through global
// memory, 4 cooperative grid barriers per iteration) this design
//   * reads S and Phi^-1 from HBM exactly once, with coalesced 8-byte loads staged through shared
//     memory, and then keeps each thread's 3*nx-entry rows of S (and Phi^-1) in REGISTERS,
//   * keeps lambda, r, p in registers and publishes p / r once per mat-vec into a small shared
//     "window" buffer (own block-rows + one halo block each side; halo of the neighbouring CTA is
//     written through DSMEM), read back with 16-byte broadcast loads,
//   * replaces both grid-wide dot products by warp shuffles + one shared-memory hop (+ one DSMEM hop
//     per CTA in cluster mode), every thread re-summing the same partials in the same order so the
//     exit decision is bit-identical everywhere (same property the reference gets from re-reducing
//     d_eta_new_temp in every block, include/pcg.cuh:189-195),
//   * fuses the axpy updates of lambda, r, p into the same pass (registers only),
//   * uses packed fp32 FMA (FFMA2, fma.rn.f32x2) for the row dot products.
// There is no grid.sync anywhere: barriers are __syncthreads (1 CTA) or barrier.cluster (cluster).
//
// Algorithm restated from reference include/pcg.cuh:118-212.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace gbdpcg {
namespace cg = cooperative_groups;

enum : int { P_NONE = 0, P_REG = 1, P_SMEM = 2, P_TMEM = 4 };   // (3 = global memory, the run-time-nx kernel's plan report)

template <typename T>
struct KernelArgs {
    const T *S;
    const T *Pinv;
    const T *gamma;
    T *lambda;
    uint32_t *iters;               // [batch] or null
    uint8_t *exits;                // [batch] or null
    unsigned long long mat_stride; // elements between consecutive problems (3*N*nx*nx)
    unsigned long long vec_stride; // elements between consecutive problems (N*nx)
    uint32_t N;                    // knot points
    uint32_t batch;
    uint32_t bpc;                  // block-rows per CTA
    uint32_t max_iter;
    uint32_t two_buffers;          // first-generation kernel (pcg_rows): 1 = S and Phi^-1 staged side by side, 0 = through one buffer in turn
    uint32_t use_tma;              // pcg_fused: 1 = matrices arrive by cp.async.bulk (16-byte aligned sources), 0 = cooperative loads
    T tol;
};

// ---------------------------------------------------------------------------------------------
// compile-time geometry
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Geo {
    static constexpr int V = 16 / (int)sizeof(T);  // elements per 16-byte shared-memory load

    // Window stride per block: a multiple of V (16-byte aligned) chosen so that the broadcast loads of
    // blocks lb and lb+2, which one warp can issue together, do not land on the same banks.
    __host__ __device__ static constexpr int nxp(int nx)
    {
        int p = (nx + V - 1) / V * V;
        return ((p * (int)sizeof(T)) % 64 == 0) ? p + V : p;
    }
    // Staging stride per block-row (elements): 3*nx*nx padded so that consecutive threads (consecutive
    // rows, running across block boundaries) hit consecutive banks when each reads element k of its row.
    __host__ __device__ static constexpr int sb(int nx)
    {
        int per = 128 / (int)sizeof(T);  // elements per bank sweep
        int base = 3 * nx * nx;
        int pad = ((nx - base) % per + per) % per;
        return base + pad;
    }
};

__host__ __device__ constexpr int round_up(int x, int m) { return (x + m - 1) / m * m; }

// Shared-memory layout (bytes), shared by host (sizing) and device (carving).
template <typename T>
struct SmemPlan {
    uint32_t win_off, red_off, cred_off, buf0_off, buf1_off, total;
};
template <typename T>
__host__ __device__ inline SmemPlan<T> make_smem_plan(int nx, int nxp, int sb, int bpc, int nbuf)
{
    SmemPlan<T> p;
    uint32_t o = 0;
    p.win_off = o;  o += round_up((bpc + 2) * nxp * (int)sizeof(T), 16);
    p.red_off = o;  o += 2 * 32 * (int)sizeof(T);   // two dot slots x up to 32 warp partials
    p.cred_off = o; o += 2 * 16 * (int)sizeof(T);   // two dot slots x up to 16 CTA partials
    p.buf0_off = o; o += (nbuf >= 1) ? round_up(bpc * sb * (int)sizeof(T), 16) : 0;
    p.buf1_off = o; o += (nbuf >= 2) ? round_up(bpc * sb * (int)sizeof(T), 16) : 0;
    p.total = o;
    (void)nx;
    return p;
}

// ---------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------
template <typename T> struct Vec16;
template <> struct Vec16<float>  { using type = float4; };
template <> struct Vec16<double> { using type = double2; };

__device__ __forceinline__ float  abs_t(float x)  { return fabsf(x); }
__device__ __forceinline__ double abs_t(double x) { return fabs(x); }

template <typename T>
__device__ __forceinline__ T warp_allsum(T x)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}

// Sum of CNT (multiple of V) shared values, fixed pairwise order, identical in every thread.
template <typename T, int CNT>
__device__ __forceinline__ T tree_sum_smem(const T *buf)
{
    constexpr int V = Geo<T>::V;
    static_assert(CNT % V == 0, "CNT must be a multiple of the vector width");
    T v[CNT];
#pragma unroll
    for (int i = 0; i < CNT / V; i++) {
        typename Vec16<T>::type q = reinterpret_cast<const typename Vec16<T>::type *>(buf)[i];
        if constexpr (V == 4) { v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w; }
        else                  { v[2 * i] = q.x; v[2 * i + 1] = q.y; }
    }
    // pairwise tree for any CNT (not only powers of two): fold the upper half onto the lower half
    int n = CNT;
#pragma unroll
    for (int step = 0; step < 6; step++) {
        if (n > 1) {
            const int half = (n + 1) / 2;
#pragma unroll
            for (int i = 0; i < CNT / 2; i++)
                if (i + half < n) v[i] += v[i + half];
            n = half;
        }
    }
    return v[0];
}

// Global -> shared staging of `nb` consecutive block-rows (3*NX*NX elements each, contiguous in global
// memory) into a buffer with block-row stride SB.  8-byte units when the source allows it (coalesced
// 256 B per warp instruction), 8 loads in flight per thread.  Padding blocks of the layout (left block
// of row 0, right block of row N-1: "never read" in the reference, include/pcg.cuh:105-106) are
// replaced by zeros so that whatever they hold (even NaN) cannot reach the arithmetic.
template <typename T, int NX>
__device__ __forceinline__ void stage_block_rows(const T *__restrict__ src, T *__restrict__ dst, int nb, int b0, int N)
{
    constexpr int BSZ = 3 * NX * NX;
    constexpr int SB = Geo<T>::sb(NX);
    constexpr int U = 8;
    const int nthr = blockDim.x, t = threadIdx.x;
    if (nb <= 0) return;

    if constexpr (sizeof(T) == 4) {
        if ((NX % 2 == 0) && ((reinterpret_cast<uintptr_t>(src) & 7u) == 0)) {
            constexpr int UPB = BSZ / 2;  // float2 units per block-row
            const int total = nb * UPB;
            const float2 *s2 = reinterpret_cast<const float2 *>(src);
            float2 *d2 = reinterpret_cast<float2 *>(dst);
            for (int base = 0; base < total; base += U * nthr) {
                float2 v[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    int q = base + u * nthr + t;
                    v[u] = make_float2(0.f, 0.f);
                    if (q < total) {
                        int lb = q / UPB, e = (q - lb * UPB) * 2, b = b0 + lb;
                        bool pad = (b == 0 && e < NX * NX) || (b == N - 1 && e >= 2 * NX * NX);
                        if (!pad) v[u] = __ldg(s2 + q);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    int q = base + u * nthr + t;
                    if (q < total) {
                        int lb = q / UPB;
                        d2[lb * (SB / 2) + (q - lb * UPB)] = v[u];
                    }
                }
            }
            return;
        }
    }
    // element-wise path: fp64 (8-byte elements already), odd nx, or a 4-byte-aligned fp32 source
    const int total = nb * BSZ;
    for (int base = 0; base < total; base += U * nthr) {
        T v[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            int q = base + u * nthr + t;
            v[u] = (T)0;
            if (q < total) {
                int lb = q / BSZ, e = q - lb * BSZ, b = b0 + lb;
                bool pad = (b == 0 && e < NX * NX) || (b == N - 1 && e >= 2 * NX * NX);
                if (!pad) v[u] = __ldg(src + q);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            int q = base + u * nthr + t;
            if (q < total) {
                int lb = q / BSZ;
                dst[lb * SB + (q - lb * BSZ)] = v[u];
            }
        }
    }
}

// Row dot product  sum_k M[k] * w[j(k)*NXP + c(k)]  over the 3*NX columns of one block-row, k = j*NX + c.
// `w` points at the window slot of block b-1.  M is read from registers (MS == nullptr) or, for a
// shared-memory-resident matrix, from MS[k*NX] (MS already points at this thread's row, column stride NX).
template <typename T, int NX, bool FROM_SMEM>
__device__ __forceinline__ T row_dot(const T (&M)[3 * NX], const T *__restrict__ MS, const T *__restrict__ w)
{
    constexpr int V = Geo<T>::V;
    constexpr int NXP = Geo<T>::nxp(NX);
    auto m = [&](int k) -> T { if constexpr (FROM_SMEM) return MS[k * NX]; else return M[k]; };

    if constexpr (sizeof(T) == 4 && NX % 2 == 0) {
        // packed fp32: two independent FFMA2 chains = four running sums
        float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < 3; j++) {
#pragma unroll
            for (int c = 0; c < NX; c += 4) {
                const float4 q = *reinterpret_cast<const float4 *>(w + j * NXP + c);
                const int k = j * NX + c;
                a0 = __ffma2_rn(make_float2(m(k), m(k + 1)), make_float2(q.x, q.y), a0);
                if (c + 2 < NX) a1 = __ffma2_rn(make_float2(m(k + 2), m(k + 3)), make_float2(q.z, q.w), a1);
            }
        }
        return (a0.x + a0.y) + (a1.x + a1.y);
    } else {
        T acc[4] = {(T)0, (T)0, (T)0, (T)0};
#pragma unroll
        for (int j = 0; j < 3; j++) {
#pragma unroll
            for (int c = 0; c < NX; c += V) {
                const typename Vec16<T>::type q = *reinterpret_cast<const typename Vec16<T>::type *>(w + j * NXP + c);
                const int k = j * NX + c;
                if constexpr (V == 4) {
                    acc[0] += m(k) * q.x;
                    if (c + 1 < NX) acc[1] += m(k + 1) * q.y;
                    if (c + 2 < NX) acc[2] += m(k + 2) * q.z;
                    if (c + 3 < NX) acc[3] += m(k + 3) * q.w;
                } else {
                    const int s = (c / 2) & 1;  // alternate between two pairs of chains
                    acc[2 * s] += m(k) * q.x;
                    if (c + 1 < NX) acc[2 * s + 1] += m(k + 1) * q.y;
                }
            }
        }
        return (acc[0] + acc[1]) + (acc[2] + acc[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <typename T, int NX, bool CLUSTERED, int PMODE, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) pcg_rows_kernel(const KernelArgs<T> a)
{
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int SB = Geo<T>::sb(NX);
    constexpr int BSZ = 3 * NX * NX;
    constexpr int NWP = round_up(round_up(MAXT, 32) / 32, Geo<T>::V);  // warp-partial slots summed per dot

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nbuf = (PMODE == P_NONE) ? 1 : (a.two_buffers ? 2 : 1);
    const SmemPlan<T> plan = make_smem_plan<T>(NX, NXP, SB, (int)a.bpc, nbuf);
    T *win = reinterpret_cast<T *>(smem_raw + plan.win_off);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);    // [2][32]
    T *cred = reinterpret_cast<T *>(smem_raw + plan.cred_off);  // [2][16]
    T *buf0 = reinterpret_cast<T *>(smem_raw + plan.buf0_off);
    T *buf1 = reinterpret_cast<T *>(smem_raw + plan.buf1_off);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int N = (int)a.N, bpc = (int)a.bpc;

    int rank = 0, csize = 1;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        rank = (int)cl.block_rank();
        csize = (int)cl.num_blocks();
    }
    const uint32_t cluster_id = blockIdx.x / (uint32_t)csize;
    const uint32_t n_clusters = gridDim.x / (uint32_t)csize;

    const int lb = t / NX, r = t - lb * NX;             // local block-row, row inside it
    const int b0 = rank * bpc;                          // first block-row of this CTA
    const int nb = max(0, min(bpc, N - b0));            // block-rows this CTA really owns
    const bool active = lb < nb;
    const int b = b0 + lb;                              // global block-row of this thread
    const int last_rank = (N - 1) / bpc;                // CTA that owns block-row N-1

    auto sync_all = [&]() {
        if constexpr (CLUSTERED) cg::this_cluster().sync();
        else __syncthreads();
    };

    // one-time init: zero window (global-boundary halos stay zero for ever) and partial-sum slots
    for (int i = t; i < (bpc + 2) * NXP; i += blockDim.x) win[i] = (T)0;
    for (int i = t; i < 64; i += blockDim.x) red[i] = (T)0;
    for (int i = t; i < 32; i += blockDim.x) cred[i] = (T)0;

    // neighbours' windows through DSMEM
    T *win_prev = nullptr, *win_next = nullptr;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        if (rank > 0) win_prev = cl.map_shared_rank(win, rank - 1);
        if (rank + 1 < csize) win_next = cl.map_shared_rank(win, rank + 1);
    }
    sync_all();  // cluster mode: also guarantees every CTA of the cluster has started before any DSMEM store

    T *const my_slot = win + (lb + 1) * NXP + r;        // where this thread publishes its vector element
    // window [x_{b-1} | x_b | x_{b+1}] of this thread's block; idle threads read slot 0 (in bounds) and
    // have their results forced to zero, so they add exactly 0 to every dot product
    const T *const my_win = win + (active ? lb : 0) * NXP;
    const bool send_prev = CLUSTERED && active && lb == 0 && rank > 0;
    const bool send_next = CLUSTERED && active && lb == nb - 1 && rank < last_rank;

    auto publish = [&](T x) {
        if (active) {
            *my_slot = x;
            if constexpr (CLUSTERED) {
                if (send_prev) win_prev[(bpc + 1) * NXP + r] = x;  // right halo of the previous CTA (always full)
                if (send_next) win_next[r] = x;                    // left halo of the next CTA
            }
        }
    };

    // all-thread dot product: returns sum over the whole system of x, bit-identical in every thread of
    // every CTA.  `slot` alternates 0/1 between consecutive dots (write-after-read safety of red/cred).
    auto all_sum = [&](T x, int slot) -> T {
        T ws = warp_allsum(x);
        T *rd = red + slot * 32;
        if (lane == 0) rd[warp] = ws;
        __syncthreads();
        T s = tree_sum_smem<T, NWP>(rd);
        if constexpr (CLUSTERED) {
            T *cr = cred + slot * 16;
            if (t < csize) {
                T *dst = cg::this_cluster().map_shared_rank(cr, t);
                dst[rank] = s;
            }
            cg::this_cluster().sync();
            s = tree_sum_smem<T, 16>(cr);
        }
        return s;
    };

    for (uint32_t prob = cluster_id; prob < a.batch; prob += n_clusters) {
        const T *Sg = a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
        const T *Pg = (PMODE != P_NONE) ? a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ : nullptr;
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + r;

        // ---- matrices: HBM -> shared (coalesced) -> registers (each thread its own row) ----
        T Sreg[3 * NX];
        T Preg[(PMODE == P_REG) ? 3 * NX : 1];
        const T *Prow = nullptr;  // P_SMEM: this thread's row inside the staging buffer
        (void)Preg; (void)Prow;

        stage_block_rows<T, NX>(Sg, buf0, nb, b0, N);
        if (PMODE != P_NONE && nbuf == 2) stage_block_rows<T, NX>(Pg, buf1, nb, b0, N);
        T lam = (T)0, gam = (T)0;
        if (active) { lam = a.lambda[vofs]; gam = __ldg(a.gamma + vofs); }
        __syncthreads();
        {
            const T *row = buf0 + lb * SB + r;
#pragma unroll
            for (int k = 0; k < 3 * NX; k++) Sreg[k] = active ? row[k * NX] : (T)0;
        }
        if constexpr (PMODE != P_NONE) {
            const T *pb = buf1;
            if (nbuf == 1) {
                __syncthreads();  // everyone has its S row; buf0 can be overwritten
                stage_block_rows<T, NX>(Pg, buf0, nb, b0, N);
                __syncthreads();
                pb = buf0;
            }
            const T *row = pb + (active ? lb * SB + r : 0);
            if constexpr (PMODE == P_REG) {
#pragma unroll
                for (int k = 0; k < 3 * NX; k++) Preg[k] = active ? row[k * NX] : (T)0;
            } else {
                Prow = row;  // P_SMEM: the rows stay where they were staged for the whole solve
            }
        }

        auto mv_S = [&]() -> T {
            const T d = row_dot<T, NX, false>(Sreg, nullptr, my_win);
            return active ? d : (T)0;
        };
        auto mv_P = [&]() -> T {
            T d = (T)0;
            if constexpr (PMODE == P_REG) d = row_dot<T, NX, false>(Preg, nullptr, my_win);
            else if constexpr (PMODE == P_SMEM) d = row_dot<T, NX, true>(Sreg, Prow, my_win);
            return active ? d : (T)0;
        };

        // ---- r = gamma - S*lambda ; r~ = Pinv r ; p = r~ ; eta = r.r~          (pcg.cuh:118-149) ----
        publish(lam);
        sync_all();
        T res = gam - mv_S();
        T rt;
        if constexpr (PMODE != P_NONE) {
            sync_all();  // window(lambda) fully consumed before it is overwritten
            publish(res);
            sync_all();
            rt = mv_P();
        } else {
            rt = res;
        }
        T p = rt;
        T eta = all_sum(res * rt, 1);

        // ---- main loop                                                          (pcg.cuh:154-208) ----
        uint32_t it = 0;
        bool hit_max = true;
        for (; it < a.max_iter; ++it) {
            publish(p);
            sync_all();
            const T ups = mv_S();                       // upsilon = S p
            const T v = all_sum(p * ups, 0);            // p . upsilon
            const T alpha = eta / v;
            lam += alpha * p;
            res -= alpha * ups;
            if constexpr (PMODE != P_NONE) {
                publish(res);
                sync_all();
                rt = mv_P();                            // r~ = Pinv r
            } else {
                rt = res;
            }
            const T eta_new = all_sum(res * rt, 1);     // r . r~
            if (abs_t(eta_new) < a.tol) { ++it; hit_max = false; break; }   // uniform: eta_new identical everywhere
            const T beta = eta_new / eta;
            eta = eta_new;
            p = rt + beta * p;
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) a.lambda[vofs] = lam;
        if (rank == 0 && t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        // Next problem: buf0/buf1/win are rewritten only after every thread has passed the last dot's
        // barrier above, and nothing reads them after that barrier.  With max_iter == 0 the init
        // barriers play that role.
        if (a.max_iter == 0) sync_all();
    }
    if constexpr (CLUSTERED) cg::this_cluster().sync();  // no CTA may exit while a neighbour can still store into it
}

}  // namespace gbdpcg
