// pcg_slab.cuh -- "column-slab" variant of the fused PCG kernel (pcg_fused.cuh): same recurrences, same
// reductions, same TMA prefetch, different ownership of the matrix entries inside a block-row.
//
// Why: in the row layout of pcg_fused.cuh a thread owns 2 rows of the nx x 3nx block-row panel and needs all
// 3*nx window values of x for them -- 168 bytes per thread per mat-vec for nx = 14, delivered by 16-byte
// shared-memory loads.  Broadcast does not make those cheaper: ncu counts 4 wavefronts per LDS.128 and the
// measured mat-vec phase (~300 clocks of a ~1400-clock iteration, twice per iteration) is exactly
// threads x 168 B / 128 B per clock.  The shared-memory -> register path, not the FMA pipe, bounds the kernel.
//
// Here the G = nx/2 threads of a block-row split the panel by COLUMNS: thread i holds all nx rows of the 6
// columns 6i..6i+5 (same 6*nx registers per matrix as before), needs only its own 6 window values (24 bytes,
// three 8-byte loads), forms nx partial row sums with nx independent FMA chains, and the G threads then
// reduce-scatter those partials with warp shuffles so that thread i ends up with the complete sums of the two
// rows it owns (2i, 2i+1).  To keep every shuffle on a compile-time register, thread i stores its matrix rows
// ROTATED by 2i: accumulator k of thread i belongs to row (2i + k) mod nx, so in round d every thread sends
// accumulators 2d, 2d+1 -- the rows owned by the thread d places further on.
// Per thread and mat-vec: 3 LDS.64 + 6*nx FFMA + 2(G-1) SHFL + 2(G-1) FADD; bytes through the shared-memory /
// shuffle crossbar: 24 + 8(G-1) = 72 for nx = 14 instead of 168.
//
// A warp holds floor(32/G) whole block-rows (groups never straddle a warp); the remaining lanes idle.
// Supported: fp32/fp64 arithmetic in T, Phi^-1 in registers or absent (P_SMEM stays with the row layout).
#pragma once
#include "pcg_fused.cuh"

namespace gbdpcg {

__host__ __device__ constexpr int slab_group(int nx) { return nx / 2; }                       // threads per block-row
__host__ __device__ constexpr int slab_blocks_per_warp(int nx) { return 32 / slab_group(nx); }
__host__ __device__ constexpr int slab_threads(int nx, int bpc)
{
    return (bpc + slab_blocks_per_warp(nx) - 1) / slab_blocks_per_warp(nx) * 32;
}

template <typename T, int NX, int PMODE, bool FUSED, bool CLUSTERED, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) pcg_slab_kernel(const KernelArgs<T> a)
{
    static_assert(NX % 2 == 0 && NX >= 2 && NX <= 64, "column-slab layout needs an even state size");
    static_assert(PMODE == P_NONE || PMODE == P_REG, "column-slab layout keeps Phi^-1 in registers");
    constexpr int G = slab_group(NX);                 // threads per block-row
    constexpr int BW = slab_blocks_per_warp(NX);      // block-rows per warp
    constexpr int CT = 6;                             // panel columns per thread: 3*NX / G
    constexpr int R = 2;                              // rows whose complete sums (and vector entries) a thread owns
    constexpr int BSZ = 3 * NX * NX;
    constexpr int NWP = round_up(round_up(MAXT, 32) / 32, 2);
    namespace cg = cooperative_groups;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = (int)a.N, bpc = (int)a.bpc;
    const bool use_tma = a.two_buffers != 0;
    constexpr int nstage = (PMODE == P_REG) ? 2 : 1;
    const FusedPlan plan = make_fused_plan<T>(NX, NX, bpc, PMODE, nstage);   // window stride = NX: x is stored gap-free
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *cred = reinterpret_cast<T *>(smem_raw + plan.cred_off);
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *stageS = reinterpret_cast<T *>(smem_raw + plan.stage_off);
    const uint32_t mbar = smem_u32(smem_raw + plan.mbar_off);
    const uint32_t mat_bytes_pad = (uint32_t)round_up(bpc * BSZ * (int)sizeof(T), 16);
    T *stageP = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(stageS) + mat_bytes_pad);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    int rank = 0, csize = 1;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        rank = (int)cl.block_rank();
        csize = (int)cl.num_blocks();
    }
    const uint32_t cluster_id = blockIdx.x / (uint32_t)csize;
    const uint32_t n_clusters = gridDim.x / (uint32_t)csize;

    const int grp = lane / G, ti = lane - grp * G;        // block-row inside the warp, thread inside the block-row
    const int lb = warp * BW + grp;                       // local block-row
    const int r0 = 2 * ti;                                // first owned row inside the block
    const int b0 = rank * bpc;
    const int nb = max(0, min(bpc, N - b0));
    const bool active = grp < BW && lb < nb;
    const int b = b0 + lb;
    const int last_rank = (N - 1) / bpc;
    const int grp_base = grp * G;                         // first lane of this thread's group

    auto sync_all = [&]() {
        if constexpr (CLUSTERED) cg::this_cluster().sync();
        else __syncthreads();
    };

    for (int i = t; i < (bpc + 2) * NX; i += blockDim.x) { winA[i] = (T)0; winB[i] = (T)0; }
    for (int i = t; i < 128; i += blockDim.x) red[i] = (T)0;
    for (int i = t; i < 64; i += blockDim.x) cred[i] = (T)0;
    if (t == 0 && use_tma) { mbar_init(mbar, 1); fence_mbar_init(); }

    T *winA_prev = nullptr, *winA_next = nullptr, *winB_prev = nullptr, *winB_next = nullptr;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        if (rank > 0) { winA_prev = cl.map_shared_rank(winA, rank - 1); winB_prev = cl.map_shared_rank(winB, rank - 1); }
        if (rank + 1 < csize) { winA_next = cl.map_shared_rank(winA, rank + 1); winB_next = cl.map_shared_rank(winB, rank + 1); }
    }
    sync_all();

    const int slot = ((active ? lb : 0) + 1) * NX + r0;   // where this thread publishes its two vector elements
    const int xofs = (active ? lb : 0) * NX + CT * ti;    // its 6 panel columns inside [x_{b-1} | x_b | x_{b+1}]
    const bool send_prev = CLUSTERED && active && lb == 0 && rank > 0;
    const bool send_next = CLUSTERED && active && lb == nb - 1 && rank < last_rank;

    auto put = [&](T *dst, const T (&x)[R]) {
        if constexpr (sizeof(T) == 4) *reinterpret_cast<float2 *>(dst) = make_float2(x[0], x[1]);
        else *reinterpret_cast<double2 *>(dst) = make_double2(x[0], x[1]);
    };
    auto publish = [&](T *win, T *win_prev, T *win_next, const T (&x)[R]) {
        if (active) {
            put(win + slot, x);
            if constexpr (CLUSTERED) {
                if (send_prev) put(win_prev + (bpc + 1) * NX + r0, x);
                if (send_next) put(win_next + r0, x);
            }
        }
    };

    int red_slot = 0;
    auto all_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 64;
        if (lane == 0) {
            if constexpr (sizeof(T) == 4) *reinterpret_cast<float2 *>(rd + 2 * warp) = make_float2(p.x(), p.y());
            else *reinterpret_cast<double2 *>(rd + 2 * warp) = make_double2(p.x(), p.y());
        }
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, NWP>(rd);
        if constexpr (CLUSTERED) {
            T *cr = cred + red_slot * 32;
            if (t < csize) {
                T *dst = cg::this_cluster().map_shared_rank(cr, t);
                dst[2 * rank] = s.x();
                dst[2 * rank + 1] = s.y();
            }
            cg::this_cluster().sync();
            s = pair_tree_sum<T, 16>(cr);
        }
        red_slot ^= 1;
        return s;
    };
    // ---- matrix prefetch (TMA) ----
    auto issue_prefetch = [&](uint32_t prob) {
        const uint32_t bytes = (uint32_t)(nb * BSZ * (int)sizeof(T));
        const T *Sg = a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
        mbar_expect_tx(mbar, (PMODE != P_NONE) ? 2 * bytes : bytes);
        bulk_g2s(smem_u32(stageS), Sg, bytes, mbar);
        if constexpr (PMODE != P_NONE) {
            const T *Pg = a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
            bulk_g2s(smem_u32(stageP), Pg, bytes, mbar);
        }
    };
    uint32_t phase = 0;
    if (use_tma && t == 0 && nb > 0 && cluster_id < a.batch) issue_prefetch(cluster_id);

    for (uint32_t prob = cluster_id; prob < a.batch; prob += n_clusters) {
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + r0;

        T lam[R], gam[R];
#pragma unroll
        for (int i = 0; i < R; i++) { lam[i] = (T)0; gam[i] = (T)0; }
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) { lam[i] = a.lambda[vofs + i]; gam[i] = __ldg(a.gamma + vofs + i); }
        }

        // ---- matrices: HBM -> shared (TMA, or cooperative loads) -> registers, rows rotated by 2*ti ----
        T Sreg[NX][CT];
        T Preg[(PMODE == P_REG) ? NX : 1][(PMODE == P_REG) ? CT : 1];
        (void)Preg;
        auto load_slab = [&](T (&M)[NX][CT], const T *img) {
            const T *base = img + (active ? lb * BSZ : 0);
#pragma unroll
            for (int c = 0; c < CT; c++) {
                const int col = CT * ti + c;
                const T *colp = base + col * NX;
#pragma unroll
                for (int kk = 0; kk < NX / 2; kk++) {
                    int row = r0 + 2 * kk;
                    row = (row >= NX) ? row - NX : row;           // accumulator k belongs to row (2*ti + k) mod NX
                    if constexpr (sizeof(T) == 4) {
                        const float2 q = *reinterpret_cast<const float2 *>(colp + row);
                        M[2 * kk][c] = q.x; M[2 * kk + 1][c] = q.y;
                    } else {
                        const double2 q = *reinterpret_cast<const double2 *>(colp + row);
                        M[2 * kk][c] = q.x; M[2 * kk + 1][c] = q.y;
                    }
                }
                // padding blocks of the layout (left of row 0, right of row N-1) are never read by the reference
                // (include/pcg.cuh:105-106): whatever they hold must not reach the arithmetic
                const bool kill = !active || (b == 0 && col < NX) || (b == N - 1 && col >= 2 * NX);
#pragma unroll
                for (int k = 0; k < NX; k++) M[k][c] = kill ? (T)0 : M[k][c];
            }
        };

        if (use_tma) {
            if (nb > 0) { mbar_wait(mbar, phase); phase ^= 1; }
        } else {
            const T *Sg = a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
            stage_image<T>(Sg, stageS, nb * BSZ);
            if constexpr (PMODE != P_NONE) {
                const T *Pg = a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
                stage_image<T>(Pg, stageP, nb * BSZ);
            }
            __syncthreads();
        }
        load_slab(Sreg, stageS);
        if constexpr (PMODE == P_REG) load_slab(Preg, stageP);
        if (use_tma) {
            __syncthreads();   // every thread holds its slab: staging may receive the next system
            const uint32_t next = prob + n_clusters;
            if (t == 0 && nb > 0 && next < a.batch) { fence_proxy_async(); issue_prefetch(next); }
        }

        // y[0..1] = complete sums of the two owned rows of  M_b . [x_{b-1} | x_b | x_{b+1}]
        auto mv = [&](const T (&M)[NX][CT], const T *win, T (&y)[R]) {
            const T *xp = win + xofs;
            T x[CT];
#pragma unroll
            for (int c = 0; c < CT; c += 2) {
                if constexpr (sizeof(T) == 4) { const float2 q = *reinterpret_cast<const float2 *>(xp + c); x[c] = q.x; x[c + 1] = q.y; }
                else { const double2 q = *reinterpret_cast<const double2 *>(xp + c); x[c] = q.x; x[c + 1] = q.y; }
            }
            T acc[NX];
#pragma unroll
            for (int k = 0; k < NX; k++) acc[k] = M[k][0] * x[0];
#pragma unroll
            for (int c = 1; c < CT; c++)
#pragma unroll
                for (int k = 0; k < NX; k++) acc[k] = fma_t(M[k][c], x[c], acc[k]);
            // reduce-scatter over the G threads of the block-row: in round d a thread hands accumulators 2d, 2d+1
            // (rows of the thread d places further on) to that thread and receives its own rows from d places back
            T y0 = acc[0], y1 = acc[1];
#pragma unroll
            for (int d = 1; d < G; d++) {
                const int src = grp_base + ((ti >= d) ? ti - d : ti - d + G);
                y0 += __shfl_sync(0xffffffffu, acc[2 * d], src);
                y1 += __shfl_sync(0xffffffffu, acc[2 * d + 1], src);
            }
            y[0] = active ? y0 : (T)0;
            y[1] = active ? y1 : (T)0;
        };
        auto mv_S = [&](const T *win, T (&y)[R]) { mv(Sreg, win, y); };
        auto mv_P = [&](const T *win, T (&y)[R]) {
            if constexpr (PMODE == P_REG) mv(Preg, win, y);
        };
        // ---- r = gamma - S*lambda                                               (pcg.cuh:118-126) ----
        // (the prologue uses whichever window the first loop publish does NOT touch, which saves a barrier)
        T res[R], rt[R], w[R], p[R];
        constexpr bool PRO_B = FUSED && PMODE != P_NONE;
        if constexpr (PRO_B) publish(winB, winB_prev, winB_next, lam);
        else publish(winA, winA_prev, winA_next, lam);
        sync_all();
        mv_S(PRO_B ? winB : winA, w);
#pragma unroll
        for (int i = 0; i < R; i++) res[i] = gam[i] - w[i];

        uint32_t it = 0;
        bool hit_max = true;

        if constexpr (FUSED) {
            // ---------------- Chronopoulos-Gear form: one reduction per iteration ----------------
            T s[R];
#pragma unroll
            for (int i = 0; i < R; i++) { p[i] = (T)0; s[i] = (T)0; }
            T rcp_gamma = (T)0, rcp_alpha = (T)0;
            for (;;) {
                if constexpr (PMODE != P_NONE) {
                    publish(winA, winA_prev, winA_next, res);
                    sync_all();
                    mv_P(winA, rt);                        // r~ = Phi^-1 r
                } else {
#pragma unroll
                    for (int i = 0; i < R; i++) rt[i] = res[i];
                }
                publish(winB, winB_prev, winB_next, rt);
                sync_all();
                mv_S(winB, w);                             // w = S r~
                T g = (T)0, d = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) { g += res[i] * rt[i]; d += w[i] * rt[i]; }
                const Pair<T> gd = all_sum(g, d);          // gamma = r.r~ , delta = w.r~
                const T gamma = gd.x(), delta = gd.y();
                if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
                if (it == a.max_iter) break;
                T beta, alpha;
                if (it == 0) { beta = (T)0; alpha = gamma / delta; }
                else {
                    beta = gamma * rcp_gamma;                                      // gamma / gamma_old
                    alpha = gamma / (delta - beta * (gamma * rcp_alpha));          // gamma / (delta - beta*gamma/alpha_old)
                }
#pragma unroll
                for (int i = 0; i < R; i++) {
                    p[i] = rt[i] + beta * p[i];
                    s[i] = w[i] + beta * s[i];
                    lam[i] += alpha * p[i];
                    res[i] -= alpha * s[i];
                }
                rcp_gamma = (T)1 / gamma;
                rcp_alpha = (T)1 / alpha;
                ++it;
            }
        } else {
            // ---------------- textbook order (reference include/pcg.cuh:131-208) ----------------
            if constexpr (PMODE != P_NONE) {
                publish(winB, winB_prev, winB_next, res);
                sync_all();
                mv_P(winB, rt);
            } else {
#pragma unroll
                for (int i = 0; i < R; i++) rt[i] = res[i];
            }
            T e0 = (T)0;
#pragma unroll
            for (int i = 0; i < R; i++) { p[i] = rt[i]; e0 += res[i] * rt[i]; }
            T eta = all_sum(e0, (T)0).x();
            for (; it < a.max_iter; ++it) {
                publish(winA, winA_prev, winA_next, p);
                sync_all();
                mv_S(winA, w);                             // upsilon = S p
                T v0 = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) v0 += p[i] * w[i];
                const T v = all_sum(v0, (T)0).x();
                const T alpha = eta / v;
#pragma unroll
                for (int i = 0; i < R; i++) { lam[i] += alpha * p[i]; res[i] -= alpha * w[i]; }
                if constexpr (PMODE != P_NONE) {
                    publish(winB, winB_prev, winB_next, res);
                    sync_all();
                    mv_P(winB, rt);                        // r~ = Phi^-1 r
                } else {
#pragma unroll
                    for (int i = 0; i < R; i++) rt[i] = res[i];
                }
                T e1 = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) e1 += res[i] * rt[i];
                const T eta_new = all_sum(e1, (T)0).x();
                if (abs_t(eta_new) < a.tol) { ++it; hit_max = false; break; }
                const T beta = eta_new / eta;
                eta = eta_new;
#pragma unroll
                for (int i = 0; i < R; i++) p[i] = rt[i] + beta * p[i];
            }
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) a.lambda[vofs + i] = lam[i];
        }
        if (rank == 0 && t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        // No barrier here: every read of this system's windows and staging happened before the barrier inside
        // the last reduction, and the partial-sum slots alternate.
    }
    if constexpr (CLUSTERED) cg::this_cluster().sync();   // no CTA may exit while a neighbour can still store into it
}

}  // namespace gbdpcg
