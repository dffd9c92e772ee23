// pcg_tmem4.cuh -- batched fp32 solver, FOUR rows per thread: two systems resident per SM, one CTA of <= 4 warps each,
// S in registers, Phi^-1 in TENSOR MEMORY.  Second generation of pcg_tmem.cuh (two rows per thread, 7-8 warps per system).
//
// Why four rows.  ncu of the two-row kernel on C5 (profiles/ncu_tmem_r1g.txt): shared-memory wavefronts 69 % of peak, FP32 pipe
// 32 %, issue 36 % -- the binding path is the window read-back (every thread loads the 3*nx window entries of its block-row
// per mat-vec, 128 B/clk/SM whether lanes broadcast or not), and its cost is  threads x 3nx x 4 B  per mat-vec, i.e. inversely
// proportional to the rows a thread owns.  With four rows a C2 system is 32 block-rows x 4 threads = 128 threads = 4 warps,
// one per tensor-memory lane quarter, and a window quad read from shared memory feeds 8 packed FMAs instead of 4.
//
// Ownership.  Block-row b is owned by G = ceil(nx/4) consecutive threads; thread g of the group owns rows 4g .. 4g+3 of the
// block-row (rows >= nx of the last thread are padding: zero matrix rows, never published).  nx = 14: 4 threads per block-row,
// the fourth with two real rows (1/8 of the FMA lanes idle; nx = 12, 16: none).  A "quad" is four consecutive columns of one of
// the three blocks of a block-row (the last quad of a block has two columns when nx = 2 mod 4); a thread's share of one matrix
// is 3*NQ quads x 4 rows.
//
// Where the matrices live (per thread): all of S (12 quads, 168 values for nx = 14) in registers except the last SQT quads,
// Phi^-1 in tensor memory except its first PQR quads.  A quad in tensor memory is one tcgen05.ld.32x32b.x16 (x8 for a half
// quad): lane = thread, so the data is thread-private; register [row*W + e] of the chunk is m(row, 4i+e).  The three on-chip
// sources are then balanced for nx = 14 (two co-resident iterations): FP32 pipe 2 x 2 x 84 FFMA2 x 2 clk = 672 clk per
// sub-partition, tensor-memory reads 2 x 168 columns x 128 lanes x 4 B / 256 B/clk = 672 clk, window reads 2 x 2 x 168 = 672
// wavefronts -- against 2 x 588 wavefronts + 2 x 420 clk of LDTM in the two-row kernel, whose two co-resident iterations take ~2200 clk.
//
// Same recurrence as pcg_tmem.cuh / pcg_fused.cuh (single-reduction Chronopoulos-Gear form of reference include/pcg.cuh:118-212),
// same exit test (pcg.cuh:195), same output semantics (pcg.cuh:212-215).  The accumulation chains differ from the two-row
// kernels (one packed chain per row, columns in ascending order), so lambda agrees with them to rounding, not bit for bit;
// parity is held against the host oracle (tests/test_parity_gpu.py::test_tmem4_*).
#pragma once
#include <type_traits>

#include "pcg_tmem.cuh"

#ifndef TMEM4_SCALAR_FMA
#define TMEM4_SCALAR_FMA 0    // 1: scalar FFMA instead of the packed FFMA2 (A/B: a packed multiply-add reads three 64-bit register operands)
#endif
#ifndef TMEM4_STAGE_PAD
#define TMEM4_STAGE_PAD 1     // staging stride per block-row padded (one bulk copy per block-row) so that the staging -> register gather is bank-conflict free
#endif
#ifndef TMEM4_L2_PREFETCH
#define TMEM4_L2_PREFETCH 1   // pull the next system's S into L2 behind the iterations (its bulk copy into the one staging buffer is exposed)
#endif

namespace gbdpcg {

#ifdef TMEM4_TRACE
// phase timeline (scripts/ubench/trace_tmem4.cu): clock64() of thread 0 of the first 8 CTAs -- per system (ordinal within the CTA):
// 0 top, 1 Phi^-1 arrived, 2 Phi^-1 in tensor memory, 3 S arrived, 4 S in registers, 5 prologue done, 6 loop left, 7 outputs written;
// per iteration of the 6th system: 0 top, 1 r~ done, 2 w done, 3 reduction done, 4 updates done
__device__ long long g_t4sys[8][40][8];
__device__ long long g_t4it[8][48][5];
#define T4S(m) do { if (t == 0 && blockIdx.x < 8 && sys_ord < 40) g_t4sys[blockIdx.x][sys_ord][m] = clock64(); } while (0)
#define T4I(m) do { if (t == 0 && blockIdx.x < 8 && sys_ord == 5 && it < 48) g_t4it[blockIdx.x][it][m] = clock64(); } while (0)
#else
#define T4S(m) do { } while (0)
#define T4I(m) do { } while (0)
#endif

__device__ __forceinline__ void l2_prefetch_bulk(const void *src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

// compile-time geometry of a thread's share
template <int NX, int SQT, int PQR>
struct T4Geo {
    static constexpr int G = (NX + 3) / 4;       // threads per block-row = quads per block
    static constexpr int NQ = G, NQT = 3 * NQ;   // quads per block, per block-row
    __host__ __device__ static constexpr int qw(int q) { return (NX - 4 * (q % NQ) >= 4) ? 4 : NX - 4 * (q % NQ); }   // columns of quad q: 4 or 2
    __host__ __device__ static constexpr bool p_tmem(int q) { return q >= PQR; }
    __host__ __device__ static constexpr bool s_tmem(int q) { return q >= NQT - SQT; }
    __host__ __device__ static constexpr bool in_tmem(bool is_s, int q) { return is_s ? s_tmem(q) : p_tmem(q); }
    __host__ __device__ static constexpr int count(bool is_s, bool tm)
    {
        int n = 0;
        for (int q = 0; q < NQT; q++) n += (in_tmem(is_s, q) == tm) ? 1 : 0;
        return n;
    }
    // k-th quad of matrix is_s that lives in tensor memory (tm) / registers (!tm)
    __host__ __device__ static constexpr int nth(bool is_s, bool tm, int k)
    {
        for (int q = 0; q < NQT; q++)
            if (in_tmem(is_s, q) == tm) {
                if (k == 0) return q;
                k--;
            }
        return -1;
    }
    // tensor-memory column of quad q: Phi^-1 quads first, then the S quads
    __host__ __device__ static constexpr int col(bool is_s, int q)
    {
        int c = 0;
        for (int k = 0; k < NQT; k++)
            if (p_tmem(k) && (is_s || k < q)) c += 4 * qw(k);
        if (is_s)
            for (int k = 0; k < q; k++)
                if (s_tmem(k)) c += 4 * qw(k);
        return c;
    }
    static constexpr int cols = col(true, NQT);
    static constexpr int alloc_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : 256;
};

// Staging -> register gather.  A thread reads, per matrix column, rows 4g..4g+1 and 4g+2..4g+3 of its block-row with two 8-byte loads;
// with the block-rows packed back to back (3nx^2 words apart) the 16 threads of a half-warp collide on the banks (nx = 14: 2.5
// wavefronts per load on average, nx = 16: 4, nx = 8: 8).  The staging image is therefore laid out in PAIRS of block-rows with
// t4_stage_pad(nx) words between pairs (one bulk copy per pair), and even block-rows read their two halves in the opposite order:
// every load is then one wavefront (layouts searched exhaustively; measured on 8192 systems: nx = 16, N = 30 0.91 -> 0.80 ms,
// nx = 8, N = 60 0.72 -> 0.64).  nx = 10, 12 (three threads per block-row) stay packed: the best layout found saves a third of the
// wavefronts and measured slower.
__host__ __device__ constexpr int t4_stage_pad(int nx)
{
#if TMEM4_STAGE_PAD
    return nx == 8 ? 8 : nx == 14 ? 24 : nx == 16 ? 16 : 0;
#else
    return 0;
#endif
}
__host__ __device__ constexpr bool t4_stage_alt(int nx) { return t4_stage_pad(nx) != 0; }
// words from the start of the staging image to block-row lb
__host__ __device__ constexpr int t4_stage_ofs(int nx, int lb) { return lb * 3 * nx * nx + (lb >> 1) * t4_stage_pad(nx); }

// bytes of one staging image (+ slack for the gather's dropped reads)
__host__ __device__ constexpr uint32_t t4_stage_bytes(int nx, int N) { return (uint32_t)((t4_stage_ofs(nx, N + 1) * 4 + 15) / 16 * 16 + 16); }

struct Tmem4Plan {
    uint32_t mbar_off, tbase_off, red_off, winA_off, winB_off, stage_off, total;
};
__host__ __device__ inline Tmem4Plan make_tmem4_plan(int nx, int nxp, int N)
{
    Tmem4Plan p;
    uint32_t o = 0;
    p.mbar_off = o;  o += 16;
    p.tbase_off = o; o += 16;
    p.red_off = o;   o += 2 * 16 * 4;                                    // 2 slots x up to 8 warps x (gamma, delta)
    p.winA_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.winB_off = o;  o += (uint32_t)round_up((N + 2) * nxp * 4, 16);
    p.stage_off = o; o += t4_stage_bytes(nx, N);                         // ONE matrix at a time (a second image follows in the solo form)
    p.total = o;
    return p;
}
constexpr int kTmem4MaxThreads = 128;    // 4 warps: one per tensor-memory lane quarter (two such CTAs per SM)
constexpr int kTmem4WideThreads = 256;   // wide form: 8 warps, two per lane quarter side by side in the columns, one CTA per SM
// Window stride per block-row (floats): exactly the four-row groups, 4 * ceil(nx/4).  (The register-row kernels pad 16-float blocks
// to 20 for their 7-threads-per-block-row broadcast pattern; here that padding made the 16-byte publish stores of two neighbouring
// block-rows collide -- 2x the wavefronts of every STS.128, 10 % of the kernel's shared-memory wavefronts, profiles/ncu_tmem4_r2d.txt.)
__host__ __device__ constexpr int t4_nxp(int nx) { return 4 * ((nx + 3) / 4); }

// ---- a thread's share of one matrix, quad by quad (compile-time recursion over the quad index) ----
// the 4 rows x W columns of quad Q from the staged block-row image (column-major nx x 3nx panel); rowp = first row of this thread
template <class GEO, int NX, int Q>
__device__ __forceinline__ void t4_gather_quad(const float *rowp, bool hi_ok, bool swap, bool kill, float (&m)[4][4])
{
    constexpr int j = Q / GEO::NQ, i = Q % GEO::NQ, W = GEO::qw(Q);
    const int o1 = swap ? 2 : 0, o2 = 2 - o1;
#pragma unroll
    for (int e = 0; e < W; e++) {
        const float *src = rowp + (j * NX + 4 * i + e) * NX;
        // even block-rows read their two halves in the opposite order (bank-conflict-free, t4_stage_pad).  Both loads are
        // unconditional, so that the lanes of a half-warp issue each of them together; a thread whose upper half does not exist
        // (rows >= nx) reads the next column's first rows (or the staging slack) and drops them.
        const float2 a = *reinterpret_cast<const float2 *>(src + o1);
        const float2 b = *reinterpret_cast<const float2 *>(src + o2);
        const float2 lo = swap ? b : a;
        float2 hi = swap ? a : b;
        hi = hi_ok ? hi : make_float2(0.f, 0.f);
        m[0][e] = kill ? 0.f : lo.x;
        m[1][e] = kill ? 0.f : lo.y;
        m[2][e] = kill ? 0.f : hi.x;
        m[3][e] = kill ? 0.f : hi.y;
    }
}
// staging -> registers (Mr[q]) or tensor memory (chunk register [row*W + e] = m(row, e)), quads Q .. NQT-1
template <class GEO, int NX, bool IS_S, int Q>
__device__ __forceinline__ void t4_load_matrix(const float *rowp, bool hi_ok, bool swap, bool kill_left, bool kill_mid, bool kill_right, uint32_t tmine,
                                               float (&Mr)[GEO::NQT][4][4])
{
    if constexpr (Q < GEO::NQT) {
        constexpr int j = Q / GEO::NQ, W = GEO::qw(Q);
        const bool kill = (j == 0) ? kill_left : (j == 2) ? kill_right : kill_mid;
        if constexpr (GEO::in_tmem(IS_S, Q)) {
            float m[4][4];
            t4_gather_quad<GEO, NX, Q>(rowp, hi_ok, swap, kill, m);
            uint32_t v[16];
#pragma unroll
            for (int row = 0; row < 4; row++)
#pragma unroll
                for (int e = 0; e < W; e++) v[row * W + e] = __float_as_uint(m[row][e]);
            if constexpr (W == 4) tmem_st16(tmine + GEO::col(IS_S, Q), v);
            else tmem_st8(tmine + GEO::col(IS_S, Q), v);
        } else {
            t4_gather_quad<GEO, NX, Q>(rowp, hi_ok, swap, kill, Mr[Q]);
        }
        t4_load_matrix<GEO, NX, IS_S, Q + 1>(rowp, hi_ok, swap, kill_left, kill_mid, kill_right, tmine, Mr);
    }
}
// window quad Q (16-byte load, 8-byte for a half quad) and its packed multiply-adds: acc[row] += m(row, 0..W-1) . x(0..W-1)
template <class GEO, int NXP, int Q>
__device__ __forceinline__ float4 t4_window_quad(const float *w)
{
    constexpr int j = Q / GEO::NQ, i = Q % GEO::NQ;
    if constexpr (GEO::qw(Q) == 4) return *reinterpret_cast<const float4 *>(w + j * NXP + 4 * i);
    else {
        const float2 h = *reinterpret_cast<const float2 *>(w + j * NXP + 4 * i);
        return make_float4(h.x, h.y, 0.f, 0.f);
    }
}
// 1 / x to within an ulp: MUFU.RCP + one Newton step, branch-free (a plain division expands into a range check with a slow-path
// call; the scalars of an iteration sit on every warp's critical path).  Operands here are dot products far from the denormal range;
// a zero or NaN operand yields inf / NaN like the division it replaces (alpha = NaN never converges: reference semantics).
__device__ __forceinline__ float t4_rcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    const float e = fmaf(-x, r, 1.f);
    return isfinite(e) ? fmaf(r, e, r) : r;
}
// acc += (m0, m1) * (x0, x1), packed (FFMA2) or as two scalar FFMAs
__device__ __forceinline__ void t4_fma2(float2 &acc, float m0, float m1, float x0, float x1)
{
#if TMEM4_SCALAR_FMA
    acc.x = fmaf(m0, x0, acc.x);
    acc.y = fmaf(m1, x1, acc.y);
#else
    acc = __ffma2_rn(make_float2(m0, m1), make_float2(x0, x1), acc);
#endif
}
template <class GEO, int NXP, int Q>
__device__ __forceinline__ void t4_fma_reg(const float (&m)[4][4], const float *w, float2 (&acc)[4])
{
    const float4 x = t4_window_quad<GEO, NXP, Q>(w);
#pragma unroll
    for (int row = 0; row < 4; row++) t4_fma2(acc[row], m[row][0], m[row][1], x.x, x.y);
    if constexpr (GEO::qw(Q) == 4) {
#pragma unroll
        for (int row = 0; row < 4; row++) t4_fma2(acc[row], m[row][2], m[row][3], x.z, x.w);
    }
}
template <class GEO, int NXP, int Q>
__device__ __forceinline__ void t4_fma_tm(const uint32_t (&u)[16], const float *w, float2 (&acc)[4])
{
    constexpr int W = GEO::qw(Q);
    const float4 x = t4_window_quad<GEO, NXP, Q>(w);
#pragma unroll
    for (int row = 0; row < 4; row++) t4_fma2(acc[row], __uint_as_float(u[row * W]), __uint_as_float(u[row * W + 1]), x.x, x.y);
    if constexpr (W == 4) {
#pragma unroll
        for (int row = 0; row < 4; row++) t4_fma2(acc[row], __uint_as_float(u[row * W + 2]), __uint_as_float(u[row * W + 3]), x.z, x.w);
    }
}
template <class GEO, bool IS_S, int K>
__device__ __forceinline__ void t4_tmem_ld(uint32_t tmine, uint32_t (&u)[16])
{
    if constexpr (K < GEO::count(IS_S, true)) {
        constexpr int q = GEO::nth(IS_S, true, K);
        if constexpr (GEO::qw(q) == 4) tmem_ld16(tmine + GEO::col(IS_S, q), u);
        else tmem_ld8(tmine + GEO::col(IS_S, q), u);
    }
}
// Step K of a mat-vec: the K-th register quad, then the K-th tensor-memory quad (wait for its chunk `cur`, start the next
// chunk into `nxt`, multiply).  The chunks are double-buffered; the caller loads chunk 0 BEFORE the barrier that publishes
// the vector, so its latency hides behind the barrier.
template <class GEO, int NXP, bool IS_S, int K>
__device__ __forceinline__ void t4_mv_step(const float (&Mr)[GEO::NQT][4][4], const float *w, uint32_t tmine, uint32_t (&cur)[16],
                                           uint32_t (&nxt)[16], float2 (&acc)[4])
{
    constexpr int nT = GEO::count(IS_S, true), nR = GEO::count(IS_S, false), nK = nT > nR ? nT : nR;
    if constexpr (K < nK) {
        if constexpr (K < nR) {
            constexpr int q = GEO::nth(IS_S, false, K);
            t4_fma_reg<GEO, NXP, q>(Mr[q], w, acc);
        }
        if constexpr (K < nT) {
            constexpr int q = GEO::nth(IS_S, true, K);
            if constexpr (GEO::qw(q) == 4) tmem_wait_ld(cur);
            else tmem_wait_ld8(cur);
            t4_tmem_ld<GEO, IS_S, K + 1>(tmine, nxt);
            t4_fma_tm<GEO, NXP, q>(cur, w, acc);
        }
        t4_mv_step<GEO, NXP, IS_S, K + 1>(Mr, w, tmine, nxt, cur, acc);
    }
}

// NWQ = warps per tensor-memory lane quarter: 1 -> systems of up to 128 four-row threads, two CTAs per SM (the C5 kernel);
// 2 -> the WIDE form, up to 256 threads (nx 14: N <= 64), one CTA per SM: warps w and w+4 keep their columns side by side.
template <int NX, int SQT, int PQR, int NWQ = 1>
__global__ void __launch_bounds__(128 * NWQ, NWQ == 1 ? 2 : 1) pcg_tmem4_kernel(const KernelArgs<float> a)
{
    static_assert(NX % 2 == 0 && NX >= 4 && NX <= 16, "even state sizes up to 16");
    static_assert(NWQ == 1 || NWQ == 2, "one or two warps per lane quarter");
    using T = float;
    using GEO = T4Geo<NX, SQT, PQR>;
    constexpr int NW = 4 * NWQ;                                                   // warps whose partials a reduction sums
    constexpr int ALLOC = NWQ == 1 ? GEO::alloc_cols : (2 * GEO::cols <= 64 ? 64 : 2 * GEO::cols <= 128 ? 128 : 2 * GEO::cols <= 256 ? 256 : 512);
    static_assert(NWQ * GEO::cols <= 512, "the warps of a lane quarter share its 512 columns");
    constexpr int R = 4, G = GEO::G, NQT = GEO::NQT, BSZ = 3 * NX * NX, NXP = t4_nxp(NX);
    static_assert(GEO::cols <= 256, "a thread's tensor-memory columns (two CTAs share the SM's 512)");
    static_assert(NXP >= 4 * G, "a thread publishes four window entries");

    extern __shared__ __align__(16) unsigned char smem_raw[];
#ifdef TMEM4_TRACE
    if (threadIdx.x == 0 && blockIdx.x < 8) g_t4it[blockIdx.x][47][0] = clock64();      // kernel entry
#endif
    const int N = (int)a.N;
    const Tmem4Plan plan = make_tmem4_plan(NX, NXP, N);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);
    uint32_t *tbase_s = reinterpret_cast<uint32_t *>(smem_raw + plan.tbase_off);
    const uint32_t mbar = smem_u32(smem_raw + plan.mbar_off);
    // Solo form (a.two_buffers: the launch has at most one CTA per SM -- single solves, batches of up to #SMs systems): S has a staging
    // image and a barrier of its own behind the first, so both matrices of a system are in flight at once and the NEXT system's S is
    // prefetched behind the iterations like its Phi^-1.  Packed form: S goes through the one image after Phi^-1 has left it.
    const bool dual = a.two_buffers != 0;
    T *stageS = dual ? reinterpret_cast<T *>(smem_raw + plan.stage_off + t4_stage_bytes(NX, N)) : stage;
    const uint32_t mbarS = dual ? mbar + 8 : mbar;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int lb = t / G, g = t - lb * G;
    const bool active = lb < N;
    const bool hi_ok = active && (4 * g + 2 < NX);         // rows 4g+2, 4g+3 exist
    const int b = lb;

    for (int i = t; i < (N + 2) * NXP; i += blockDim.x) { winA[i] = 0.f; winB[i] = 0.f; }
    if (t < 4 * NW) red[t] = 0.f;
    if (t == 0) { mbar_init(mbar, 1); mbar_init(mbar + 8, 1); fence_mbar_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tbase_s)), "r"(ALLOC) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *tbase_s;
    const uint32_t tmine = tbase + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * GEO::cols);

    const int slot = (lb + 1) * NXP + 4 * g;
    const int wofs = (active ? lb : 0) * NXP;
    auto publish = [&](T *win, const T (&x)[R]) {
        if (active) *reinterpret_cast<float4 *>(win + slot) = make_float4(x[0], x[1], x[2], x[3]);   // padding rows hold 0
    };
    int red_slot = 0;
    auto all_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 2 * NW;
        if (lane == 0) *reinterpret_cast<float2 *>(rd + 2 * warp) = make_float2(p.x(), p.y());
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, NW>(rd);
        red_slot ^= 1;
        return s;
    };

    const uint32_t mat_bytes = (uint32_t)(N * BSZ * (int)sizeof(T));
    // matrix image -> staging: one bulk copy (packed staging, thread 0), or one per pair of block-rows spread over the warps (padded
    // staging: a bulk copy is issued lane by lane, so the pairs are dealt to the warps first).  Called by every thread.
    auto issue_copy = [&](const T *src, T *dst, const uint32_t bar) {
        if constexpr (t4_stage_pad(NX) == 0) {
            if (t == 0) {
                fence_proxy_async();
                mbar_expect_tx(bar, mat_bytes);
                bulk_g2s(smem_u32(dst), src, mat_bytes, bar);
            }
        } else {
            const int nwarps = (int)blockDim.x >> 5, npairs = (N + 1) >> 1;
            const int c = warp + nwarps * lane;                       // pair of block-rows this lane copies
            if (t == 0) mbar_expect_tx(bar, mat_bytes);
            if (c < npairs) {
                fence_proxy_async();
                const int rows = (2 * c + 2 <= N) ? 2 : 1;
                bulk_g2s(smem_u32(dst + t4_stage_ofs(NX, 2 * c)), src + (size_t)(2 * c) * BSZ, (uint32_t)(rows * BSZ * sizeof(T)), bar);
            }
        }
    };
    // one barrier: Phi^-1 completes its even phases, S the odd ones; two barriers: each alternates on its own
    uint32_t nP = 0, nS = 0;
    if (blockIdx.x < a.batch) {
        issue_copy(a.Pinv + (size_t)blockIdx.x * a.mat_stride, stage, mbar);
        if (dual) issue_copy(a.S + (size_t)blockIdx.x * a.mat_stride, stageS, mbarS);
#if TMEM4_L2_PREFETCH
        // packed form: the first system's S follows its Phi^-1 through the one staging image; start its HBM read now, so that the two
        // cold reads of a launch's first wave overlap instead of queueing (first-system "wait S" 3283 -> L2-hit time)
        else if (t == 32) l2_prefetch_bulk(a.S + (size_t)blockIdx.x * a.mat_stride, mat_bytes);
#endif
    }

    const bool kill_left = !active || b == 0, kill_right = !active || b == N - 1, kill_mid = !active;
    const int rofs = active ? t4_stage_ofs(NX, lb) + 4 * g : 0;
    const T *rowp = stage + rofs, *rowpS = stageS + rofs;
    const bool swap = t4_stage_alt(NX) && !(lb & 1);

    int sys_ord = -1;
    (void)sys_ord;
    for (uint32_t prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
        ++sys_ord;
        T4S(0);
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + 4 * g;
        T lam[R] = {0.f, 0.f, 0.f, 0.f}, gam[R] = {0.f, 0.f, 0.f, 0.f};
        if (active) {
            lam[0] = a.lambda[vofs]; lam[1] = a.lambda[vofs + 1];
            gam[0] = __ldg(a.gamma + vofs); gam[1] = __ldg(a.gamma + vofs + 1);
            if (hi_ok) {
                lam[2] = a.lambda[vofs + 2]; lam[3] = a.lambda[vofs + 3];
                gam[2] = __ldg(a.gamma + vofs + 2); gam[3] = __ldg(a.gamma + vofs + 3);
            }
        }

        // ---- Phi^-1: staging -> tensor memory (registers for its first PQR quads) ----
        T Pr[NQT][4][4], Sr[NQT][4][4];               // only the quads that stay in registers are ever touched
        mbar_wait(mbar, dual ? (nP & 1u) : 0u); ++nP;
        T4S(1);
        t4_load_matrix<GEO, NX, false, 0>(rowp, hi_ok, swap, kill_left, kill_mid, kill_right, tmine, Pr);
        __syncthreads();                                   // every thread has read its Phi^-1 rows: staging is free
        T4S(2);
        {
            const uint32_t next = prob + gridDim.x;
            if (!dual) issue_copy(a.S + (size_t)prob * a.mat_stride, stage, mbar);
            else if (next < a.batch) issue_copy(a.Pinv + (size_t)next * a.mat_stride, stage, mbar);
        }

        // ---- S: staging -> registers (tensor memory for its last SQT quads) ----
        mbar_wait(mbarS, dual ? (nS & 1u) : 1u); ++nS;
        T4S(3);
        t4_load_matrix<GEO, NX, true, 0>(rowpS, hi_ok, swap, kill_left, kill_mid, kill_right, tmine, Sr);
        tmem_wait_st();
        __syncthreads();                                   // staging is free again: fetch the next system's Phi^-1 behind the iterations
        T4S(4);
        {
            const uint32_t next = prob + gridDim.x;
            if (next < a.batch) {
                if (dual) issue_copy(a.S + (size_t)next * a.mat_stride, stageS, mbarS);
                else {
                    issue_copy(a.Pinv + (size_t)next * a.mat_stride, stage, mbar);
#if TMEM4_L2_PREFETCH
                    if (t == 32) l2_prefetch_bulk(a.S + (size_t)next * a.mat_stride, mat_bytes);
#endif
                }
            }
        }

        // y = M * window for this thread's four rows: quads in registers and quads in tensor memory alternate (t4_mv_step)
        auto finish = [&](const float2 (&acc)[R], T (&y)[R]) {
            y[0] = active ? acc[0].x + acc[0].y : 0.f;
            y[1] = active ? acc[1].x + acc[1].y : 0.f;
            y[2] = hi_ok ? acc[2].x + acc[2].y : 0.f;
            y[3] = hi_ok ? acc[3].x + acc[3].y : 0.f;
        };
        auto mv_S = [&](const T *win, uint32_t (&ua)[16], T (&y)[R]) {
            uint32_t ub[16];
            float2 acc[R] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
            t4_mv_step<GEO, NXP, true, 0>(Sr, win + wofs, tmine, ua, ub, acc);
            finish(acc, y);
        };
        auto mv_P = [&](const T *win, uint32_t (&ua)[16], T (&y)[R]) {
            uint32_t ub[16];
            float2 acc[R] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
            t4_mv_step<GEO, NXP, false, 0>(Pr, win + wofs, tmine, ua, ub, acc);
            finish(acc, y);
        };

        // ---- r = gamma - S*lambda                                               (pcg.cuh:118-126) ----
        T res[R], rt[R], w[R], p[R], s[R];
        uint32_t u0[16];
        publish(winB, lam);
        t4_tmem_ld<GEO, true, 0>(tmine, u0);
        __syncthreads();
        mv_S(winB, u0, w);
#pragma unroll
        for (int i = 0; i < R; i++) { res[i] = gam[i] - w[i]; p[i] = 0.f; s[i] = 0.f; }

        uint32_t it = 0;
        bool hit_max = true;
        T rcp_gamma = 0.f, rcp_alpha = 0.f;
        T4S(5);
        for (;;) {
            T4I(0);
            publish(winA, res);
            t4_tmem_ld<GEO, false, 0>(tmine, u0);
            __syncthreads();
            mv_P(winA, u0, rt);                   // r~ = Phi^-1 r                (pcg.cuh:131-140, 181-184)
            T4I(1);
            publish(winB, rt);
            t4_tmem_ld<GEO, true, 0>(tmine, u0);
            __syncthreads();
            mv_S(winB, u0, w);                       // w = S r~
            T4I(2);
            T gg = 0.f, dd = 0.f;
#pragma unroll
            for (int i = 0; i < R; i++) { gg += res[i] * rt[i]; dd += w[i] * rt[i]; }
            const Pair<T> gd = all_sum(gg, dd);            // gamma = r.r~ , delta = w.r~
            const T gamma = gd.x(), delta = gd.y();
            T4I(3);
            if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
            if (it == a.max_iter) break;
            // beta = gamma / gamma_old, alpha = gamma / (delta - beta * gamma / alpha_old).  Both divisions of an iteration are
            // reciprocals of numbers known right after the reduction (gamma and the denominator), so they are issued together;
            // 1 / alpha for the next iteration is denominator / gamma, no third division.
            const T rg = t4_rcp(gamma);
            const T beta = (it == 0) ? 0.f : gamma * rcp_gamma;
            const T denom = (it == 0) ? delta : delta - beta * (gamma * rcp_alpha);
            const T alpha = gamma * t4_rcp(denom);
#pragma unroll
            for (int i = 0; i < R; i++) {
                p[i] = rt[i] + beta * p[i];
                lam[i] += alpha * p[i];
                s[i] = w[i] + beta * s[i];
                res[i] -= alpha * s[i];
            }
            rcp_gamma = rg;
            rcp_alpha = denom * rg;
            T4I(4);
            ++it;
        }

        T4S(6);
        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
            a.lambda[vofs] = lam[0]; a.lambda[vofs + 1] = lam[1];
            if (hi_ok) { a.lambda[vofs + 2] = lam[2]; a.lambda[vofs + 3] = lam[3]; }
        }
        if (t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        T4S(7);
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(ALLOC) : "memory");
#ifdef TMEM4_TRACE
    if (threadIdx.x == 0 && blockIdx.x < 8) g_t4it[blockIdx.x][47][1] = clock64();      // kernel end
#endif
}

}  // namespace gbdpcg
