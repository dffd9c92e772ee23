// pcg_tmem.cuh -- batched fp32 solver with TWO (small systems: FOUR) systems resident per SM: Phi^-1 and the right blocks of S live in
// TENSOR MEMORY (256 KB per SM that the register-row kernel of pcg_fused.cuh leaves idle), the left and diagonal
// blocks of S in registers.
//
// Why: one C2 system (nx 14, N 32) fills the register file (S + Phi^-1 rows = 168 of 255 registers x 224 threads), so
// pcg_fused runs ONE system per SM and every iteration is a chain of dependent phases (mat-vec, shuffle tree, barrier,
// scalars) with 7 warps to cover it: issue slots 26 % busy, 0.16 of the HBM line.  scripts/ubench/tmem_rate.cu measured
// the missing piece: tcgen05.ld.32x32b streams thread-private data TMEM -> registers at >= 272 B/clk per SM next to,
// and not instead of, the 128 B/clk of shared-memory loads the window reads already use.  With two thirds of a system's
// matrix entries in TMEM a thread needs 4nx matrix registers instead of 12nx, two CTAs fit on an SM (<= 128 registers,
// <= 113 KB shared memory, 256 TMEM columns each), and one system's reduction / barrier latencies are covered by the
// other's mat-vec.
//
// Same recurrence (single-reduction Chronopoulos-Gear form, reference include/pcg.cuh:104-217 restated as in
// pcg_fused.cuh), same ownership (a thread owns 2 rows), same accumulation chains and the same reduction trees as
// pcg_fused_kernel<float, NX, 2, P_REG, 1, 0>, so lambda and the iteration counts are BIT-IDENTICAL to that kernel
// (tests/test_parity_gpu.py::test_tmem_*).
//
// TMEM layout.  Warp w may touch TMEM lanes 32*(w%4) .. +31 only; lane = the thread's lane, so data is thread-private.
// Columns (32-bit) per thread: 128 = 8 chunks of 16; warps w and w+4 share a lane quarter and take columns 0..127 and
// 128..255 of the CTA's 256-column allocation (a CTA of at most 4 warps allocates 128 columns only, so four such CTAs share an
// SM).  Chunk (j, cb) of a matrix holds, for window block j in {left, diag,
// right} and quad pair cb, the entries m(row, j*nx + 8*cb + 4*q + e) at register 8*q + 4*row + e -- exactly the operands
// of two 16-byte window loads, in the order the FFMA2 chains consume them.  Phi^-1: chunks 0 .. 3*CPB-1, S right block:
// the following CPB chunks (CPB = chunks per block = ceil(ceil(nx/4)/2) <= 2, i.e. nx <= 16).
#pragma once
#include "pcg_fused.cuh"

namespace gbdpcg {

struct TmemPlan {
    uint32_t mbar_off, tbase_off, red_off, winA_off, winB_off, stage_off, total;
};
__host__ __device__ inline TmemPlan make_tmem_plan(int nx, int nxp, int N)
{
    TmemPlan p;
    uint32_t o = 0;
    p.mbar_off = o;  o += 8;
    p.tbase_off = o; o += 8;
    p.red_off = o;   o += 2 * 64 * 4;
    p.winA_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.winB_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.stage_off = o; o += (uint32_t)round_up(N * 3 * nx * nx * 4, 16);   // ONE matrix at a time
    p.total = o;
    return p;
}
// TMEM columns per CTA: up to 4 warps each own a lane quarter and 128 columns are enough (four CTAs per SM); 5..8 warps share
// the quarters two by two and take 256 (two CTAs per SM fill the 512 columns).
__host__ __device__ constexpr int tmem_cols(int threads) { return threads <= 128 ? 128 : 256; }
constexpr int kTmemMaxThreads = 256;    // 8 warps: two per TMEM lane quarter

template <int NX>
__global__ void __launch_bounds__(kTmemMaxThreads, 2) pcg_tmem_kernel(const KernelArgs<float> a)
{
    static_assert(NX % 2 == 0 && NX >= 2 && NX <= 16, "TMEM layout: even state sizes up to 16");
    using T = float;
    constexpr int R = 2, G = NX / 2, BSZ = 3 * NX * NX, NXP = Geo<T>::nxp(NX);
    constexpr int NQ = (NX + 3) / 4;       // 16-byte window loads per block
    constexpr int CPB = (NQ + 1) / 2;      // TMEM chunks per block
    constexpr int NCP = 3 * CPB;           // chunks of Phi^-1
    constexpr int NCH = FUSED_NCH_R2;      // same accumulation chains as block_row_mv
    constexpr int NWP = 8;
    static_assert((NCP + CPB) * 16 <= 128, "a thread's TMEM columns");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = (int)a.N;
    const TmemPlan plan = make_tmem_plan(NX, NXP, N);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);
    uint32_t *tbase_s = reinterpret_cast<uint32_t *>(smem_raw + plan.tbase_off);
    const uint32_t mbar = smem_u32(smem_raw + plan.mbar_off);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int lb = t / G, r0 = R * (t - lb * G);
    const bool active = lb < N;
    const int b = lb;

    for (int i = t; i < (N + 2) * NXP; i += blockDim.x) { winA[i] = 0.f; winB[i] = 0.f; }
    for (int i = t; i < 128; i += blockDim.x) red[i] = 0.f;
    if (t == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tbase_s)), "r"(tmem_cols((int)blockDim.x)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *tbase_s;
    const uint32_t tmine = tbase + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(128 * (warp >> 2));
    const uint32_t tP = tmine, tS = tmine + 16 * NCP;

    const int slot = (lb + 1) * NXP + r0;
    const int wofs = (active ? lb : 0) * NXP;
    auto publish = [&](T *win, const T (&x)[R]) {
        if (active) *reinterpret_cast<float2 *>(win + slot) = make_float2(x[0], x[1]);
    };
    int red_slot = 0;
    auto all_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 64;
        if (lane == 0) *reinterpret_cast<float2 *>(rd + 2 * warp) = make_float2(p.x(), p.y());
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, NWP>(rd);
        red_slot ^= 1;
        return s;
    };

    const uint32_t mat_bytes = (uint32_t)(N * BSZ * (int)sizeof(T));
    auto issue_copy = [&](const T *src) {
        mbar_expect_tx(mbar, mat_bytes);
        bulk_g2s(smem_u32(stage), src, mat_bytes, mbar);
    };
    uint32_t phase = 0;
    if (t == 0 && blockIdx.x < a.batch) issue_copy(a.Pinv + (size_t)blockIdx.x * a.mat_stride);

    const bool kill_left = !active || b == 0, kill_right = !active || b == N - 1, kill_mid = !active;
    // 16 entries of chunk (j, cb) of this thread's two rows, from the staged block-row image (column-major panel)
    auto gather_chunk = [&](const T *row, int j, int cb, uint32_t (&v)[16]) {
        const bool kill = (j == 0) ? kill_left : (j == 2) ? kill_right : kill_mid;
#pragma unroll
        for (int q = 0; q < 2; q++)
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int c = 8 * cb + 4 * q + e;      // column inside the block
                float2 m = make_float2(0.f, 0.f);
                if (c < NX) m = *reinterpret_cast<const float2 *>(row + (j * NX + c) * NX);
                v[8 * q + e] = kill ? 0u : __float_as_uint(m.x);
                v[8 * q + 4 + e] = kill ? 0u : __float_as_uint(m.y);
            }
    };
    // window quad i of a block: the last quad of a state size that is 2 mod 4 holds only two entries -- an 8-byte load costs
    // 2 shared-memory wavefronts instead of 4.  Written out for the reader; nvcc narrows the unused half by itself
    // (C5 0.996 vs 0.998 ms with and without).
    auto load_quad = [&](const T *src, int i) -> float4 {
        if (4 * i + 2 >= NX) {
            const float2 h = *reinterpret_cast<const float2 *>(src);
            return make_float4(h.x, h.y, 0.f, 0.f);
        }
        return *reinterpret_cast<const float4 *>(src);
    };
    // the FFMA2 chains of block_row_mv, restricted to the two window quads of one chunk
    auto fma_chunk = [&](int j, int cb, const uint32_t (&u)[16], const T *w, float2 (&acc)[R][NCH]) {
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int i = 2 * cb + q;
            if (i < NQ) {
                const float4 x = load_quad(w + j * NXP + 4 * i, i);
                const int c = 4 * i;
                const int pair = j * ((NX + 1) / 2) + 2 * i;
#pragma unroll
                for (int row = 0; row < R; row++) {
                    const int o = 8 * q + 4 * row;
                    acc[row][pair % NCH] = __ffma2_rn(make_float2(__uint_as_float(u[o]), __uint_as_float(u[o + 1])),
                                                      make_float2(x.x, x.y), acc[row][pair % NCH]);
                    if (c + 2 < NX)
                        acc[row][(pair + 1) % NCH] = __ffma2_rn(make_float2(__uint_as_float(u[o + 2]), __uint_as_float(u[o + 3])),
                                                                make_float2(x.z, x.w), acc[row][(pair + 1) % NCH]);
                }
            }
        }
    };
    auto finish = [&](float2 (&acc)[R][NCH], T (&y)[R]) {
#pragma unroll
        for (int row = 0; row < R; row++) {
            float2 s = acc[row][0];
#pragma unroll
            for (int ch = 1; ch < NCH; ch++) s = __fadd2_rn(s, acc[row][ch]);
            y[row] = active ? s.x + s.y : 0.f;
        }
    };

    for (uint32_t prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + r0;
        T lam[R] = {0.f, 0.f}, gam[R] = {0.f, 0.f};
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) { lam[i] = a.lambda[vofs + i]; gam[i] = __ldg(a.gamma + vofs + i); }
        }
        const T *row = stage + (active ? lb * BSZ + r0 : 0);

        // ---- Phi^-1: staging -> tensor memory ----
        mbar_wait(mbar, phase); phase ^= 1;
#pragma unroll
        for (int c = 0; c < NCP; c++) {
            uint32_t v[16];
            gather_chunk(row, c / CPB, c % CPB, v);
            tmem_st16(tP + 16 * c, v);
        }
        __syncthreads();                                   // every thread has read its Phi^-1 rows: staging is free
        if (t == 0) { fence_proxy_async(); issue_copy(a.S + (size_t)prob * a.mat_stride); }

        // ---- S: left and diagonal blocks -> registers, right block -> tensor memory ----
        mbar_wait(mbar, phase); phase ^= 1;
        T Sreg[R][2 * NX];
#pragma unroll
        for (int k = 0; k < 2 * NX; k++) {
            const float2 q = *reinterpret_cast<const float2 *>(row + k * NX);
            const bool kill = (k < NX) ? kill_left : kill_mid;
            Sreg[0][k] = kill ? 0.f : q.x;
            Sreg[1][k] = kill ? 0.f : q.y;
        }
#pragma unroll
        for (int c = 0; c < CPB; c++) {
            uint32_t v[16];
            gather_chunk(row, 2, c, v);
            tmem_st16(tS + 16 * c, v);
        }
        tmem_wait_st();
        __syncthreads();                                   // staging is free again: fetch the next system's Phi^-1 behind the iterations
        {
            const uint32_t next = prob + gridDim.x;
            if (t == 0 && next < a.batch) { fence_proxy_async(); issue_copy(a.Pinv + (size_t)next * a.mat_stride); }
        }

        // The first tensor-memory load of a mat-vec does not depend on the window: the caller issues it BEFORE the barrier
        // that publishes the vector (u0), so its latency hides behind the barrier.
        auto mv_S = [&](const T *win, uint32_t (&ua)[16], T (&y)[R]) {
            const T *w = win + wofs;
            uint32_t ub[16];
            float2 acc[R][NCH];
#pragma unroll
            for (int row_ = 0; row_ < R; row_++)
#pragma unroll
                for (int ch = 0; ch < NCH; ch++) acc[row_][ch] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 2; j++)
#pragma unroll
                for (int i = 0; i < NQ; i++) {
                    const float4 x = load_quad(w + j * NXP + 4 * i, i);
                    const int c = 4 * i, k = j * NX + c;
                    const int pair = j * ((NX + 1) / 2) + 2 * i;
#pragma unroll
                    for (int row_ = 0; row_ < R; row_++) {
                        acc[row_][pair % NCH] =
                            __ffma2_rn(make_float2(Sreg[row_][k], Sreg[row_][k + 1]), make_float2(x.x, x.y), acc[row_][pair % NCH]);
                        if (c + 2 < NX)
                            acc[row_][(pair + 1) % NCH] = __ffma2_rn(make_float2(Sreg[row_][k + 2], Sreg[row_][k + 3]),
                                                                     make_float2(x.z, x.w), acc[row_][(pair + 1) % NCH]);
                    }
                }
            tmem_wait_ld(ua);
            if constexpr (CPB == 2) tmem_ld16(tS + 16, ub);
            fma_chunk(2, 0, ua, w, acc);
            if constexpr (CPB == 2) {
                tmem_wait_ld(ub);
                fma_chunk(2, 1, ub, w, acc);
            }
            finish(acc, y);
        };
        auto mv_P = [&](const T *win, uint32_t (&u0)[16], T (&y)[R]) {
            const T *w = win + wofs;
            uint32_t u1[16];
            float2 acc[R][NCH];
#pragma unroll
            for (int row_ = 0; row_ < R; row_++)
#pragma unroll
                for (int ch = 0; ch < NCH; ch++) acc[row_][ch] = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < NCP; c++) {
                if (c & 1) {
                    tmem_wait_ld(u1);
                    if (c + 1 < NCP) tmem_ld16(tP + 16 * (c + 1), u0);
                    fma_chunk(c / CPB, c % CPB, u1, w, acc);
                } else {
                    tmem_wait_ld(u0);
                    if (c + 1 < NCP) tmem_ld16(tP + 16 * (c + 1), u1);
                    fma_chunk(c / CPB, c % CPB, u0, w, acc);
                }
            }
            finish(acc, y);
        };

        // ---- r = gamma - S*lambda                                               (pcg.cuh:118-126) ----
        T res[R], rt[R], w[R], p[R], s[R];
        uint32_t u0[16];
        publish(winB, lam);
        tmem_ld16(tS, u0);
        __syncthreads();
        mv_S(winB, u0, w);
#pragma unroll
        for (int i = 0; i < R; i++) { res[i] = gam[i] - w[i]; p[i] = 0.f; s[i] = 0.f; }

        uint32_t it = 0;
        bool hit_max = true;
        T rcp_gamma = 0.f, rcp_alpha = 0.f;
        for (;;) {
            publish(winA, res);
            tmem_ld16(tP, u0);
            __syncthreads();
            mv_P(winA, u0, rt);                            // r~ = Phi^-1 r                (pcg.cuh:131-140, 181-184)
            publish(winB, rt);
            tmem_ld16(tS, u0);
            __syncthreads();
            mv_S(winB, u0, w);                             // w = S r~
            T g = 0.f, d = 0.f;
#pragma unroll
            for (int i = 0; i < R; i++) { g += res[i] * rt[i]; d += w[i] * rt[i]; }
            const Pair<T> gd = all_sum(g, d);              // gamma = r.r~ , delta = w.r~
            const T gamma = gd.x(), delta = gd.y();
            if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
            if (it == a.max_iter) break;
            T beta, alpha;
            if (it == 0) { beta = 0.f; alpha = gamma / delta; }
            else {
                beta = gamma * rcp_gamma;
                alpha = gamma / (delta - beta * (gamma * rcp_alpha));
            }
#pragma unroll
            for (int i = 0; i < R; i++) {
                p[i] = rt[i] + beta * p[i];
                lam[i] += alpha * p[i];
                s[i] = w[i] + beta * s[i];
                res[i] -= alpha * s[i];
            }
            rcp_gamma = 1.f / gamma;
            rcp_alpha = 1.f / alpha;
            ++it;
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) a.lambda[vofs + i] = lam[i];
        }
        if (t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(tmem_cols((int)blockDim.x)) : "memory");
}

}  // namespace gbdpcg
