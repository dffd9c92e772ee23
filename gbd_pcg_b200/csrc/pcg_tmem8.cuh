// pcg_tmem8.cuh -- batched fp32 solver, four rows per thread PAIR: the columns of a row group are split between two neighbouring
// lanes, so a system of up to 128 row groups runs on 8 warps and an SM holds 16 warps of two systems.
//
// Why.  pcg_tmem4.cuh (one thread per group of four rows, 4 warps per system, two systems per SM) is latency bound, not pipe bound:
// ncu of C5 (profiles/ncu_tmem4_r2b.txt) has the FP32 pipe 41 % busy, shared-memory wavefronts at 34 % of peak, issue slots 38 %,
// and the phase timeline (profiles/trace_tmem4_r2c.txt) shows an iteration of one CTA taking ~1500 clocks of which the two mat-vecs
// are 750 and the reduction + scalars + barrier skew the rest -- with ONE warp per sub-partition per system nothing hides those
// latencies except the other system's single warp.  Splitting a thread's 3nx columns between two lanes halves every warp's
// mat-vec chain and doubles the warps (4 per sub-partition) without changing what is read from shared memory: a thread loads only
// the window quads of its own columns.
//
// Ownership.  Thread t: pair p = t / 2, half h = t % 2; block-row lb = p / G, row group g = p % G (rows 4g .. 4g+3), G = nx / 4
// rounded up and EVEN (nx = 8, 14, 16).  Of the 3*NQ quads (four columns each; columns >= nx of a block are zero padding) half h
// owns quads 2k + h, k = 0 .. NOWN-1.  After a mat-vec the two lanes exchange partial sums with one shuffle each way and lane h keeps
// rows 4g+2h, 4g+2h+1: lambda, r, p, s live once, the dot products count every row once.
// Per thread: Phi^-1 quads in tensor memory (NOWN x 16 columns), the first NOWN-SQT quads of S in registers, the last SQT in tensor
// memory; at most 128 columns per thread (warps w and w+4 share a lane quarter; two CTAs share the SM's 512 columns).
//
// Same recurrence, exit test and output semantics as pcg_tmem4.cuh (reference include/pcg.cuh:118-212, :195, :212-215).
#pragma once
#include "pcg_tmem4.cuh"

namespace gbdpcg {

template <int NX, int SQT>
struct T8Geo {
    static constexpr int G = (NX + 3) / 4, NQ = G, NQT = 3 * NQ, NOWN = NQT / 2;
    static_assert(NQ % 2 == 0, "the two quads of a step lie in the same block");
    static constexpr int SREG = NOWN - SQT;
    static constexpr int p_cols = 16 * NOWN, cols = 16 * (NOWN + SQT);
    static_assert(cols <= 128, "a thread's tensor-memory columns");
    // window / panel position of step k for half 0 (half 1: the next quad of the same block)
    __host__ __device__ static constexpr int blk(int k) { return (2 * k) / NQ; }
    __host__ __device__ static constexpr int quad0(int k) { return (2 * k) % NQ; }
};

struct Tmem8Plan {
    uint32_t mbar_off, tbase_off, red_off, winA_off, winB_off, stage_off, total;
};
__host__ __device__ inline Tmem8Plan make_tmem8_plan(int nx, int nxp, int N)
{
    Tmem8Plan p;
    uint32_t o = 0;
    p.mbar_off = o;  o += 16;
    p.tbase_off = o; o += 16;
    p.red_off = o;   o += 2 * 16 * 4;                                    // 2 slots x 8 warps x (gamma, delta)
    p.winA_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.winB_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.stage_off = o; o += (uint32_t)round_up(t4_stage_ofs(nx, N + 1) * 4, 16) + 64;
    p.total = o;
    return p;
}
constexpr int kTmem8MaxThreads = 256;

template <int NX, int SQT>
__global__ void __launch_bounds__(kTmem8MaxThreads, 2) pcg_tmem8_kernel(const KernelArgs<float> a)
{
    using T = float;
    using GEO = T8Geo<NX, SQT>;
    constexpr int G = GEO::G, NQ = GEO::NQ, NOWN = GEO::NOWN, SREG = GEO::SREG, BSZ = 3 * NX * NX, NXP = t4_nxp(NX);
    static_assert(NXP >= 4 * G, "window padding covers the zero columns of the last quad");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = (int)a.N;
    const Tmem8Plan plan = make_tmem8_plan(NX, NXP, N);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);
    uint32_t *tbase_s = reinterpret_cast<uint32_t *>(smem_raw + plan.tbase_off);
    const uint32_t mbar = smem_u32(smem_raw + plan.mbar_off);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int pr = t >> 1, h = t & 1;
    const int lb = pr / G, g = pr - lb * G;
    const bool active = lb < N;
    const bool hi_ok = active && (4 * g + 2 < NX);              // rows 4g+2, 4g+3 exist
    const bool own_ok = active && (4 * g + 2 * h < NX);         // the two rows this lane keeps (4g+2h, 4g+2h+1) exist
    const int b = lb;
    const int tcols = ((int)blockDim.x > 128) ? 256 : 128;      // warps w and w+4 share a lane quarter: 128 columns each

    for (int i = t; i < (N + 2) * NXP; i += blockDim.x) { winA[i] = 0.f; winB[i] = 0.f; }
    if (t < 32) red[t] = 0.f;
    if (t == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tbase_s)), "r"(tcols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *tbase_s;
    const uint32_t tmine = tbase + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(128 * (warp >> 2));
    const uint32_t tP = tmine, tS = tmine + GEO::p_cols;

    const int slot = (lb + 1) * NXP + 4 * g + 2 * h;
    const int wofs = (active ? lb : 0) * NXP + 4 * h;
    auto publish = [&](T *win, const T (&x)[2]) {
        if (own_ok) *reinterpret_cast<float2 *>(win + slot) = make_float2(x[0], x[1]);
    };
    int red_slot = 0;
    auto all_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 16;
        if (lane == 0) *reinterpret_cast<float2 *>(rd + 2 * warp) = make_float2(p.x(), p.y());
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, 8>(rd);
        red_slot ^= 1;
        return s;
    };

    const uint32_t mat_bytes = (uint32_t)(N * BSZ * (int)sizeof(T));
    auto issue_copy = [&](const T *src) {                       // called by every thread (see pcg_tmem4.cuh)
        if constexpr (t4_stage_pad(NX) == 0) {
            if (t == 0) {
                fence_proxy_async();
                mbar_expect_tx(mbar, mat_bytes);
                bulk_g2s(smem_u32(stage), src, mat_bytes, mbar);
            }
        } else {
            const int nwarps = (int)blockDim.x >> 5, npairs = (N + 1) >> 1;
            const int c = warp + nwarps * lane;
            if (t == 0) mbar_expect_tx(mbar, mat_bytes);
            if (c < npairs) {
                fence_proxy_async();
                const int rows = (2 * c + 2 <= N) ? 2 : 1;
                bulk_g2s(smem_u32(stage + t4_stage_ofs(NX, 2 * c)), src + (size_t)(2 * c) * BSZ, (uint32_t)(rows * BSZ * sizeof(T)), mbar);
            }
        }
    };
    uint32_t phase = 0;
    if (blockIdx.x < a.batch) issue_copy(a.Pinv + (size_t)blockIdx.x * a.mat_stride);

    const bool kill_left = !active || b == 0, kill_right = !active || b == N - 1, kill_mid = !active;
    // first of this thread's columns in the staged block-row image (column-major nx x 3nx panel): rows 4g.., columns of quad h
    const T *rowp = stage + (active ? t4_stage_ofs(NX, lb) + 4 * g + 4 * h * NX : 0);
    // 4 rows x 4 columns of this thread's quad of step k; columns >= nx of the block and rows >= nx are zero
    auto gather_quad = [&](int k, T (&m)[4][4]) {
        const int j = GEO::blk(k), i0 = GEO::quad0(k);
        const bool kill = (j == 0) ? kill_left : (j == 2) ? kill_right : kill_mid;
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const bool col_ok = 4 * (i0 + h) + e < NX;
            const T *src = rowp + (j * NX + 4 * i0 + (col_ok ? e : 0)) * NX;
            const float2 lo = *reinterpret_cast<const float2 *>(src);
            float2 hi = *reinterpret_cast<const float2 *>(src + 2);
            hi = hi_ok ? hi : make_float2(0.f, 0.f);
            const bool z = kill || !col_ok;
            m[0][e] = z ? 0.f : lo.x;
            m[1][e] = z ? 0.f : lo.y;
            m[2][e] = z ? 0.f : hi.x;
            m[3][e] = z ? 0.f : hi.y;
        }
    };
    auto to_chunk = [&](const T (&m)[4][4], uint32_t (&v)[16]) {
#pragma unroll
        for (int row = 0; row < 4; row++)
#pragma unroll
            for (int e = 0; e < 4; e++) v[row * 4 + e] = __float_as_uint(m[row][e]);
    };
    // acc[row] += m(row, 0..3) . x(0..3)
    auto fma_regs = [&](const T (&m)[4][4], const float4 x, float2 (&acc)[4]) {
#pragma unroll
        for (int row = 0; row < 4; row++) acc[row] = __ffma2_rn(make_float2(m[row][0], m[row][1]), make_float2(x.x, x.y), acc[row]);
#pragma unroll
        for (int row = 0; row < 4; row++) acc[row] = __ffma2_rn(make_float2(m[row][2], m[row][3]), make_float2(x.z, x.w), acc[row]);
    };
    auto fma_chunk = [&](const uint32_t (&u)[16], const float4 x, float2 (&acc)[4]) {
#pragma unroll
        for (int row = 0; row < 4; row++)
            acc[row] = __ffma2_rn(make_float2(__uint_as_float(u[row * 4]), __uint_as_float(u[row * 4 + 1])), make_float2(x.x, x.y), acc[row]);
#pragma unroll
        for (int row = 0; row < 4; row++)
            acc[row] = __ffma2_rn(make_float2(__uint_as_float(u[row * 4 + 2]), __uint_as_float(u[row * 4 + 3])), make_float2(x.z, x.w), acc[row]);
    };
    // the two lanes of a pair hold partial sums over their own columns: lane h keeps rows 2h, 2h+1 and receives the partner's part
    auto combine = [&](const float2 (&acc)[4], T (&y)[2]) {
        const T y0 = acc[0].x + acc[0].y, y1 = acc[1].x + acc[1].y, y2 = acc[2].x + acc[2].y, y3 = acc[3].x + acc[3].y;
        const T keep0 = h ? y2 : y0, keep1 = h ? y3 : y1, give0 = h ? y0 : y2, give1 = h ? y1 : y3;
        const T got0 = __shfl_xor_sync(0xffffffffu, give0, 1), got1 = __shfl_xor_sync(0xffffffffu, give1, 1);
        // half 0 + half 1 in both lanes' order of summation (the sum of two floats commutes)
        y[0] = own_ok ? keep0 + got0 : 0.f;
        y[1] = own_ok ? keep1 + got1 : 0.f;
    };

    for (uint32_t prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + 4 * g + 2 * h;
        T lam[2] = {0.f, 0.f}, gam[2] = {0.f, 0.f};
        if (own_ok) {
            lam[0] = a.lambda[vofs]; lam[1] = a.lambda[vofs + 1];
            gam[0] = __ldg(a.gamma + vofs); gam[1] = __ldg(a.gamma + vofs + 1);
        }

        // ---- Phi^-1: staging -> tensor memory ----
        mbar_wait(mbar, phase); phase ^= 1;
#pragma unroll
        for (int k = 0; k < NOWN; k++) {
            T m[4][4];
            uint32_t v[16];
            gather_quad(k, m);
            to_chunk(m, v);
            tmem_st16(tP + 16 * k, v);
        }
        __syncthreads();                                   // every thread has read its Phi^-1 columns: staging is free
        issue_copy(a.S + (size_t)prob * a.mat_stride);

        // ---- S: staging -> registers (first SREG quads) / tensor memory (the rest) ----
        mbar_wait(mbar, phase); phase ^= 1;
        T Sr[SREG > 0 ? SREG : 1][4][4];
#pragma unroll
        for (int k = 0; k < NOWN; k++) {
            if (k < SREG) gather_quad(k, Sr[k]);
            else {
                T m[4][4];
                uint32_t v[16];
                gather_quad(k, m);
                to_chunk(m, v);
                tmem_st16(tS + 16 * (k - SREG), v);
            }
        }
        tmem_wait_st();
        __syncthreads();                                   // staging is free again: the next system's Phi^-1 arrives behind the iterations
        {
            const uint32_t next = prob + gridDim.x;
            if (next < a.batch) {
                issue_copy(a.Pinv + (size_t)next * a.mat_stride);
                if (t == 32) l2_prefetch_bulk(a.S + (size_t)next * a.mat_stride, mat_bytes);
            }
        }

        auto window_quad = [&](const T *win, int k) -> float4 {
            return *reinterpret_cast<const float4 *>(win + wofs + GEO::blk(k) * NXP + 4 * GEO::quad0(k));
        };
        // Phi^-1 x: all quads from tensor memory, double-buffered; the caller loaded chunk 0 before the barrier
        auto mv_P = [&](const T *win, uint32_t (&ua)[16], T (&y)[2]) {
            uint32_t ub[16];
            float2 acc[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int k = 0; k < NOWN; k++) {
                const float4 x = window_quad(win, k);
                if (k & 1) {
                    tmem_wait_ld(ub);
                    if (k + 1 < NOWN) tmem_ld16(tP + 16 * (k + 1), ua);
                    fma_chunk(ub, x, acc);
                } else {
                    tmem_wait_ld(ua);
                    if (k + 1 < NOWN) tmem_ld16(tP + 16 * (k + 1), ub);
                    fma_chunk(ua, x, acc);
                }
            }
            combine(acc, y);
        };
        // S x: register quads first (the tensor-memory chunk 0 of S is in flight meanwhile), then the tensor-memory quads
        auto mv_S = [&](const T *win, uint32_t (&ua)[16], T (&y)[2]) {
            uint32_t ub[16];
            float2 acc[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int k = 0; k < SREG; k++) fma_regs(Sr[k], window_quad(win, k), acc);
#pragma unroll
            for (int k = SREG; k < NOWN; k++) {
                const float4 x = window_quad(win, k);
                if ((k - SREG) & 1) {
                    tmem_wait_ld(ub);
                    if (k + 1 < NOWN) tmem_ld16(tS + 16 * (k + 1 - SREG), ua);
                    fma_chunk(ub, x, acc);
                } else {
                    tmem_wait_ld(ua);
                    if (k + 1 < NOWN) tmem_ld16(tS + 16 * (k + 1 - SREG), ub);
                    fma_chunk(ua, x, acc);
                }
            }
            combine(acc, y);
        };

        // ---- r = gamma - S*lambda                                               (pcg.cuh:118-126) ----
        T res[2], rt[2], w[2], p[2], s[2];
        uint32_t u0[16];
        publish(winB, lam);
        if constexpr (SQT > 0) tmem_ld16(tS, u0);
        __syncthreads();
        mv_S(winB, u0, w);
#pragma unroll
        for (int i = 0; i < 2; i++) { res[i] = gam[i] - w[i]; p[i] = 0.f; s[i] = 0.f; }

        uint32_t it = 0;
        bool hit_max = true;
        T rcp_gamma = 0.f, rcp_alpha = 0.f;
        for (;;) {
            publish(winA, res);
            tmem_ld16(tP, u0);
            __syncthreads();
            mv_P(winA, u0, rt);                            // r~ = Phi^-1 r                (pcg.cuh:131-140, 181-184)
            publish(winB, rt);
            if constexpr (SQT > 0) tmem_ld16(tS, u0);
            __syncthreads();
            mv_S(winB, u0, w);                             // w = S r~
            const T gg = res[0] * rt[0] + res[1] * rt[1], dd = w[0] * rt[0] + w[1] * rt[1];
            const Pair<T> gd = all_sum(gg, dd);            // gamma = r.r~ , delta = w.r~
            const T gamma = gd.x(), delta = gd.y();
            if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
            if (it == a.max_iter) break;
            const T rg = t4_rcp(gamma);
            const T beta = (it == 0) ? 0.f : gamma * rcp_gamma;
            const T denom = (it == 0) ? delta : delta - beta * (gamma * rcp_alpha);
            const T alpha = gamma * t4_rcp(denom);
#pragma unroll
            for (int i = 0; i < 2; i++) {
                p[i] = rt[i] + beta * p[i];
                lam[i] += alpha * p[i];
                s[i] = w[i] + beta * s[i];
                res[i] -= alpha * s[i];
            }
            rcp_gamma = rg;
            rcp_alpha = denom * rg;
            ++it;
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (own_ok) { a.lambda[vofs] = lam[0]; a.lambda[vofs + 1] = lam[1]; }
        if (t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(tcols) : "memory");
}

}  // namespace gbdpcg
