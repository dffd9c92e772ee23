// pcg_fused.cuh -- the sm_100a PCG kernel (second generation of this library's hot path).
//
// One thread-block cluster (1..16 CTAs) owns one system for the whole solve and is persistent over a
// batch of systems.  A thread owns R rows (R = 2 for fp32, 1 for fp64; 2 in the fp64 P_TMEM form) of the (N*nx)-dimensional system
// and keeps those rows of S and Phi^-1 in REGISTERS; lambda, r, p, ... live in registers too.  The only
// data that moves per iteration is the vector being multiplied: it is published into a small shared
// "window" (own block-rows + one halo block on each side; the halo of the neighbouring CTA is written
// through DSMEM) and read back with 16-byte broadcast loads.
//
// What the reference does per iteration (include/pcg.cuh:154-208): 2 block-row mat-vecs through shared
// memory with nx of 64 threads working, 2 grid-wide dot products (global partials + grid.sync +
// redundant re-reduction), 4 grid.sync.  What this kernel does per iteration:
//   FUSED = true  (the default): Chronopoulos-Gear form of the same recurrence -- r~ = Phi^-1 r and
//       w = S r~ are formed first, then gamma = r.r~ and delta = w.r~ are reduced TOGETHER, and
//       beta = gamma/gamma_old, alpha = gamma/(delta - beta*gamma/alpha_old), p = r~ + beta p,
//       s = w + beta s (= S p), lambda += alpha p, r -= alpha s.  Same iterates in exact arithmetic, same
//       exit test on the same quantity (|r.Phi^-1 r| < tol after each update), ONE reduction and three
//       block barriers per iteration.
//   FUSED = false (GBDPCG_FLAG_TEXTBOOK): the reference's own order of operations
//       (two reductions, four barriers).
// Reductions are warp shuffles + one shared-memory hop (+ one DSMEM hop per CTA in cluster mode); every
// thread re-sums the same partials in the same order, so alpha, beta and the exit decision are
// bit-identical in every thread of every CTA (the reference gets this from re-reducing d_eta_new_temp in
// every block, include/pcg.cuh:189-195).  There is no grid.sync and no global-memory traffic inside the loop.
//
// Where Phi^-1 lives (PMODE): P_REG registers (the default where both matrices fit), P_SMEM resident in shared memory, P_NONE identity,
// P_TMEM tensor memory -- the fp64 throughput form: two rows per thread, the rows of Phi^-1 and the tail of the rows of S in TMEM, half
// the SMs per system (TmGeo / tm_* below; instantiated in pcg_fused_tmem_inst.cu).
//
// Batched mode: while a system is being solved, the NEXT system's S and Phi^-1 block-rows are prefetched
// into shared memory by the TMA engine (cp.async.bulk + mbarrier), so HBM latency is hidden behind the
// iterations and every matrix byte is read from HBM exactly once.
//
// Algorithm restated from reference include/pcg.cuh:118-212.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "pcg_rows.cuh"  // KernelArgs, Geo, Vec16, abs_t, warp helpers, stage_block_rows, P_* enums
#include "tmem_ops.cuh"  // tcgen05.ld / st wrappers (P_TMEM)

#ifndef FUSED_NCH_R2
#define FUSED_NCH_R2 2
#endif
#ifndef FUSED_LOAD64
#define FUSED_LOAD64 1
#endif
#ifndef FUSED_QBUF
#define FUSED_QBUF 2
#endif
#ifndef FUSED_NCH_F64
#define FUSED_NCH_F64 4   // accumulation chains per row of the fp64 mat-vec.  14 chains shorten the mat-vec in the trace harness
                          // (713 -> 519 clk) but cost spills in the real kernels: C3 55.3 -> 57.4 us, (10,5,20) 36.9 -> 38.9 us
#endif
#ifndef FUSED_GENERIC_HOIST
#define FUSED_GENERIC_HOIST 1   // generic (fp64 / odd-nx) mat-vec: issue a window block's loads together, ahead of its multiply-adds
#endif
#ifndef FUSED_WAIT_CLUSTER_SCOPE
// 1: the receiver waits with try_wait.acquire.cluster (what the memory model asks for data pushed by another CTA).  0: the
// default CTA-scope form, which spares the CCTL.IVALL ptxas emits after a cluster-scope acquire -- measured: no difference
// (36.80 vs 36.83 us on the C4 trace), so the formally correct form stays.
#define FUSED_WAIT_CLUSTER_SCOPE 1
#endif
#ifndef FUSED_GHOST
#define FUSED_GHOST 1     // cluster mode, single-reduction recurrence: neighbours' boundary block-rows of r are kept as ghost
                          // copies (updated locally from w, alpha, beta) instead of being exchanged -> one cluster barrier less
#endif
#ifndef FUSED_PUSH
#define FUSED_PUSH 1      // cluster mode with ghost copies: halo rows and reduction partials travel as st.async pushes that
                          // complete a transaction count on the RECEIVER's mbarrier; no barrier.cluster inside the loop
#endif
#ifndef FUSED_SCALAR_FMA
#define FUSED_SCALAR_FMA 0   // 1: fp32 mat-vec with scalar FFMA (A/B against the packed FFMA2 form)
#endif

namespace gbdpcg {

// ---------------------------------------------------------------------------------------------
// shared-memory plan (host sizes it, device carves it)
// ---------------------------------------------------------------------------------------------
struct FusedPlan {
    uint32_t mbar_off, red_off, cred_off, winA_off, winB_off, stage_off, psm_off, ghost_off, xbar_off, tbase_off, total;
};
// nstage: number of matrices the staging area holds side by side (2: S and Phi^-1 for P_REG; else 1)
// PMODE == P_SMEM keeps Phi^-1 resident in its own area (psm) for the whole solve.
template <typename T>
__host__ __device__ inline FusedPlan make_fused_plan(int nx, int nxp, int bpc, int pmode, int nstage)
{
    FusedPlan p;
    uint32_t o = 0;
    p.mbar_off = o;  o += 16;
    p.red_off = o;   o += 2 * 2 * 32 * (uint32_t)sizeof(T);       // 2 slots x [warp][2] partials, up to 32 warps
    p.cred_off = o;  o += 2 * 2 * 16 * (uint32_t)sizeof(T);       // 2 slots x [cta][2] partials, up to 16 CTAs
    p.winA_off = o;  o += (uint32_t)round_up((bpc + 2) * nxp * (int)sizeof(T), 16);
    p.winB_off = o;  o += (uint32_t)round_up((bpc + 2) * nxp * (int)sizeof(T), 16);
    const uint32_t mat = (uint32_t)round_up(bpc * 3 * nx * nx * (int)sizeof(T), 16);
    p.psm_off = o;   o += (pmode == P_SMEM) ? mat : 0;
    p.stage_off = o; o += (uint32_t)nstage * mat;
    // cluster-mode extras go last, so that the single-CTA layout (the batched hot path) is what it was measured with
    p.ghost_off = o; o += (uint32_t)round_up(8 * nxp * (int)sizeof(T), 16);   // ghost r [2 sides], s [2], w [2 parities][2 sides]
    p.xbar_off = o;  o += 32;                                                 // push mode: halo[2] and reduction[2] mbarriers
    p.tbase_off = o; o += (pmode == P_TMEM) ? 16 : 0;                         // tensor-memory base address (tcgen05.alloc writes it)
    p.total = o;
    return p;
}

// ---------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + bulk async copy (TMA engine, 1-D)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ---- one-sided pushes between the CTAs of a cluster: the store lands in the receiver's shared memory and completes
// `bytes` of the transaction count of the receiver's mbarrier, which the receiver alone waits on ----
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t cta_rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
    return r;
}
__device__ __forceinline__ void st_async(uint32_t raddr, float a, uint32_t rmbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f32 [%0], %1, [%2];" ::"r"(raddr), "f"(a), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void st_async(uint32_t raddr, double a, uint32_t rmbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(raddr), "d"(a), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void st_async2(uint32_t raddr, float a, float b, uint32_t rmbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(raddr), "f"(a), "f"(b),
                 "r"(rmbar)
                 : "memory");
}
__device__ __forceinline__ void st_async2(uint32_t raddr, double a, double b, uint32_t rmbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];" ::"r"(raddr), "d"(a), "d"(b),
                 "r"(rmbar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t mbar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAITC_%=:\n"
#if FUSED_WAIT_CLUSTER_SCOPE
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
#else
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
#endif
        "@p bra DONEC_%=;\n"
        "bra WAITC_%=;\n"
        "DONEC_%=:\n"
        "}\n" ::"r"(mbar), "r"(parity) : "memory");
}

// Cooperative global -> shared copy of `count` contiguous elements (the unpadded image of this CTA's block-rows).
// Used when the TMA path is not available (source not 16-byte aligned, or Phi^-1 resident in shared memory).
template <typename T>
__device__ __noinline__ void stage_image(const T *__restrict__ src, T *__restrict__ dst, int count)
{
    constexpr int U = 8;
    const int nthr = blockDim.x, t = threadIdx.x;
    if (count <= 0) return;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && (count * (int)sizeof(T)) % 16 == 0) {
        const int total = count * (int)sizeof(T) / 16;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4 *d4 = reinterpret_cast<uint4 *>(dst);
        for (int base = 0; base < total; base += U * nthr) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < total) v[u] = __ldg(s4 + q); }
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < total) d4[q] = v[u]; }
        }
    } else {
        for (int base = 0; base < count; base += U * nthr) {
            T v[U];
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < count) v[u] = __ldg(src + q); }
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < count) dst[q] = v[u]; }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// packed arithmetic: a "pair" is two T values that are reduced / updated together
// ---------------------------------------------------------------------------------------------
template <typename T> struct Pair;
template <> struct Pair<float> {
    float2 v;
    __device__ __forceinline__ Pair() {}
    __device__ __forceinline__ Pair(float a, float b) : v(make_float2(a, b)) {}
    __device__ __forceinline__ float x() const { return v.x; }
    __device__ __forceinline__ float y() const { return v.y; }
    __device__ __forceinline__ Pair operator+(const Pair &o) const { Pair r; r.v = __fadd2_rn(v, o.v); return r; }
};
template <> struct Pair<double> {
    double a, b;
    __device__ __forceinline__ Pair() {}
    __device__ __forceinline__ Pair(double a_, double b_) : a(a_), b(b_) {}
    __device__ __forceinline__ double x() const { return a; }
    __device__ __forceinline__ double y() const { return b; }
    __device__ __forceinline__ Pair operator+(const Pair &o) const { return Pair(a + o.a, b + o.b); }
};
template <typename T>
__device__ __forceinline__ Pair<T> shfl_xor_pair(const Pair<T> &p, int m)
{
    return Pair<T>(__shfl_xor_sync(0xffffffffu, p.x(), m), __shfl_xor_sync(0xffffffffu, p.y(), m));
}

// Sum of CNT (value0, value1) pairs stored interleaved in shared memory; fixed pairwise order, identical
// in every thread.  CNT is a multiple of 2 (fp32: one 16-byte load = 2 pairs) or any (fp64: 1 pair per load).
template <typename T, int CNT>
__device__ __forceinline__ Pair<T> pair_tree_sum(const T *buf)
{
    Pair<T> v[CNT];
    if constexpr (sizeof(T) == 4) {
        static_assert(CNT % 2 == 0, "");
#pragma unroll
        for (int i = 0; i < CNT / 2; i++) {
            const float4 q = reinterpret_cast<const float4 *>(buf)[i];
            v[2 * i] = Pair<T>(q.x, q.y);
            v[2 * i + 1] = Pair<T>(q.z, q.w);
        }
    } else {
#pragma unroll
        for (int i = 0; i < CNT; i++) {
            const double2 q = reinterpret_cast<const double2 *>(buf)[i];
            v[i] = Pair<T>(q.x, q.y);
        }
    }
    int n = CNT;
#pragma unroll
    for (int step = 0; step < 6; step++) {
        if (n > 1) {
            const int half = (n + 1) / 2;
#pragma unroll
            for (int i = 0; i < CNT / 2; i++)
                if (i + half < n) v[i] = v[i] + v[i + half];
            n = half;
        }
    }
    return v[0];
}

// ---------------------------------------------------------------------------------------------
// block-row mat-vec for the R rows a thread owns:  y[row] = sum_k M[row][k] * w[j(k)*NXP + c(k)],  k = j*NX + c
// `w` points at the window slot of block b-1.  M comes from registers, or (SMEM_M) from shared memory
// where element (row, k) sits at Ms[k*NX + row] (the global column-major image of the block-row).
// ---------------------------------------------------------------------------------------------
template <typename T, int NX, int R, bool SMEM_M>
__device__ __forceinline__ void block_row_mv(const T (&M)[R][3 * NX], const T *__restrict__ Ms, const T *__restrict__ w, T (&y)[R])
{
    constexpr int NXP = Geo<T>::nxp(NX);
    auto m = [&](int row, int k) -> T {
        if constexpr (SMEM_M) return Ms[k * NX + row];
        else return M[row][k];
    };
    if constexpr (sizeof(T) == 4 && NX % 2 == 0 && !FUSED_SCALAR_FMA) {
        constexpr int NCH = (R == 1) ? 4 : FUSED_NCH_R2;  // independent FFMA2 chains per row
        constexpr int NQ = (NX + 3) / 4;       // 16-byte loads per window block
        float2 acc[R][NCH];
#pragma unroll
        for (int row = 0; row < R; row++)
#pragma unroll
            for (int ch = 0; ch < NCH; ch++) acc[row][ch] = make_float2(0.f, 0.f);
        // Software pipeline over the three window blocks with at most two blocks of loads in flight: with all
        // twelve 16-byte loads hoisted the register file (2 x 3*NX matrix entries per row) overflows.
        float4 q[2][NQ];
        auto load_block = [&](int j, float4 (&dst)[NQ]) {
#pragma unroll
            for (int i = 0; i < NQ; i++) dst[i] = *reinterpret_cast<const float4 *>(w + j * NXP + 4 * i);
        };
        auto fma_block = [&](int j, const float4 (&src)[NQ]) {
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const int c = 4 * i, k = j * NX + c;
                const int pair = j * ((NX + 1) / 2) + 2 * i;
#pragma unroll
                for (int row = 0; row < R; row++) {
                    acc[row][pair % NCH] =
                        __ffma2_rn(make_float2(m(row, k), m(row, k + 1)), make_float2(src[i].x, src[i].y), acc[row][pair % NCH]);
                    if (c + 2 < NX)
                        acc[row][(pair + 1) % NCH] = __ffma2_rn(make_float2(m(row, k + 2), m(row, k + 3)),
                                                                make_float2(src[i].z, src[i].w), acc[row][(pair + 1) % NCH]);
                }
            }
        };
#if FUSED_QBUF == 2
        load_block(0, q[0]);
        load_block(1, q[1]);
        asm volatile("" ::: "memory");
        fma_block(0, q[0]);
        load_block(2, q[0]);
        asm volatile("" ::: "memory");
        fma_block(1, q[1]);
        fma_block(2, q[0]);
#else
        load_block(0, q[0]);
        asm volatile("" ::: "memory");
        fma_block(0, q[0]);
        asm volatile("" ::: "memory");
        load_block(1, q[0]);
        asm volatile("" ::: "memory");
        fma_block(1, q[0]);
        asm volatile("" ::: "memory");
        load_block(2, q[0]);
        asm volatile("" ::: "memory");
        fma_block(2, q[0]);
#endif
#pragma unroll
        for (int row = 0; row < R; row++) {
            float2 s = acc[row][0];
#pragma unroll
            for (int ch = 1; ch < NCH; ch++) s = __fadd2_rn(s, acc[row][ch]);
            y[row] = s.x + s.y;
        }
    } else {
        constexpr int V = Geo<T>::V;
        // independent accumulation chains per row (knob for fp64, see FUSED_NCH_F64)
        constexpr int NCH = (sizeof(T) == 8) ? FUSED_NCH_F64 : 4;
        T acc[R][NCH];
#pragma unroll
        for (int row = 0; row < R; row++)
#pragma unroll
            for (int ch = 0; ch < NCH; ch++) acc[row][ch] = (T)0;
        int idx = 0;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            // all 16-byte loads of one window block first, then its multiply-adds: with one or two warps per sub-partition
            // nothing else hides a load-use stall, and ptxas otherwise sinks every load next to its first use
            constexpr int NQ = (NX + V - 1) / V;
            typename Vec16<T>::type qv[NQ];
#pragma unroll
            for (int i = 0; i < NQ; i++) qv[i] = *reinterpret_cast<const typename Vec16<T>::type *>(w + j * NXP + i * V);
            if constexpr (FUSED_GENERIC_HOIST != 0) asm volatile("" ::: "memory");
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const int c = i * V, k = j * NX + c;
                T qq[4];
                if constexpr (V == 4) { qq[0] = qv[i].x; qq[1] = qv[i].y; qq[2] = qv[i].z; qq[3] = qv[i].w; }
                else { qq[0] = qv[i].x; qq[1] = qv[i].y; qq[2] = (T)0; qq[3] = (T)0; }
#pragma unroll
                for (int e = 0; e < V; e++) {
                    if (c + e < NX) {
#pragma unroll
                        for (int row = 0; row < R; row++) acc[row][(idx + e) % NCH] += m(row, k + e) * qq[e];
                    }
                }
                idx += V;
            }
            if constexpr (FUSED_GENERIC_HOIST != 0) asm volatile("" ::: "memory");
        }
#pragma unroll
        for (int row = 0; row < R; row++) {
#pragma unroll
            for (int n = NCH; n > 1; n = (n + 1) / 2) {
#pragma unroll
                for (int i = 0; i < n / 2; i++) acc[row][i] = acc[row][i] + acc[row][i + (n + 1) / 2];
            }
            y[row] = acc[row][0];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// PMODE == P_TMEM (fp64 throughput mode).  The R rows of Phi^-1 a thread owns -- and the last SK panel columns of its rows
// of S -- live in TENSOR MEMORY instead of registers, so a thread can own R = 2 fp64 rows (S alone would fill its 255
// registers twice over together with Phi^-1) and a system needs half the SMs: C3 (nx 14, N 64, fp64, 602 KB of matrices)
// runs on a 2-CTA cluster of 224 threads where the register-row plan needs 4 CTAs.
// Layout, per thread (lane = the thread's lane, so the data is thread-private): panel column k of the block-row (k = j*NX + c,
// j = left / diag / right block) takes R * (sizeof(T)/4) consecutive 32-bit columns, rows interleaved, so that one 16-column
// tcgen05.ld delivers KPC consecutive panel columns of all R rows -- the operands of KPC/V 16-byte window loads.
// Columns [0, PC): Phi^-1;  [PC, PC + SK*CPK): S, panel columns K0 .. 3NX-1.  Warps w and w+4 share a lane quarter and take
// consecutive column ranges of CPT columns each.
// ---------------------------------------------------------------------------------------------
template <typename T, int NX, int R, int SK>
struct TmGeo {
    static constexpr int W32 = (int)sizeof(T) / 4;      // 32-bit columns per element
    static constexpr int CPK = R * W32;                 // columns per panel column
    static constexpr int KPC = 16 / CPK;                // panel columns per 16-column chunk
    static constexpr int K0 = 3 * NX - SK;              // S: panel columns below K0 stay in registers
    static constexpr int PC = 3 * NX * CPK;             // columns of Phi^-1
    static constexpr int CPT = PC + SK * CPK;           // columns per thread
    static_assert(16 % CPK == 0 && KPC >= 2, "a chunk holds whole panel columns");
    static_assert(SK >= 0 && SK <= 3 * NX && SK % KPC == 0, "the S tail is whole chunks");
};
__host__ __device__ constexpr int tm_alloc_cols(int cpt, int threads)
{
    const int need = ((threads + 127) / 128) * cpt;     // warps per lane quarter x columns per thread
    return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512;
}
__device__ __forceinline__ float fma_t(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return __fma_rn(a, b, c); }
// element <-> 32-bit tensor-memory words (one for float, two for double), and element e of a 16-byte window vector
__device__ __forceinline__ void tm_pack(uint32_t *v, float x) { v[0] = __float_as_uint(x); }
__device__ __forceinline__ void tm_pack(uint32_t *v, double x) { v[0] = (uint32_t)__double2loint(x); v[1] = (uint32_t)__double2hiint(x); }
__device__ __forceinline__ void tm_unpack(const uint32_t *v, float &x) { x = __uint_as_float(v[0]); }
__device__ __forceinline__ void tm_unpack(const uint32_t *v, double &x) { x = __hiloint2double((int)v[1], (int)v[0]); }
__device__ __forceinline__ float tm_vec_get(const float4 &v, int e) { return e == 0 ? v.x : e == 1 ? v.y : e == 2 ? v.z : v.w; }
__device__ __forceinline__ float tm_vec_get(const float2 &v, int e) { return e == 0 ? v.x : v.y; }
// window loads of the P_TMEM mat-vec: 16 bytes where a block is a whole number of them, else 8 (fp32 with nx = 2 mod 4: nx 10, 14)
template <typename T, int NX> struct TmWin { static constexpr int V = Geo<T>::V; using Vec = typename Vec16<T>::type; };
template <> struct TmWin<float, 10> { static constexpr int V = 2; using Vec = float2; };
template <> struct TmWin<float, 14> { static constexpr int V = 2; using Vec = float2; };
template <> struct TmWin<float, 6> { static constexpr int V = 2; using Vec = float2; };
__device__ __forceinline__ double tm_vec_get(const double2 &v, int e) { return e == 0 ? v.x : v.y; }

// rows M[row][k], k in [KB, KE)  ->  tensor-memory columns from taddr on
template <typename T, int NX, int R, int KB, int KE>
__device__ __forceinline__ void tm_store_rows(const T (&M)[R][3 * NX], uint32_t taddr)
{
    constexpr int W32 = (int)sizeof(T) / 4, CPK = R * W32, KPC = 16 / CPK, NK = KE - KB, NFULL = NK / KPC, REM = NK % KPC;
    static_assert(REM == 0 || REM * CPK == 8, "the last chunk is whole or a half");
#pragma unroll
    for (int c = 0; c < NFULL; c++) {
        uint32_t v[16];
#pragma unroll
        for (int kk = 0; kk < KPC; kk++)
#pragma unroll
            for (int row = 0; row < R; row++) tm_pack(&v[(kk * R + row) * W32], M[row][KB + c * KPC + kk]);
        tmem_st16(taddr + 16 * c, v);
    }
    if constexpr (REM != 0) {
        uint32_t v[16];
#pragma unroll
        for (int kk = 0; kk < REM; kk++)
#pragma unroll
            for (int row = 0; row < R; row++) tm_pack(&v[(kk * R + row) * W32], M[row][KB + NFULL * KPC + kk]);
        tmem_st8(taddr + 16 * NFULL, v);
    }
}

// acc[row][k % NCH] += m(row, k) * x(k) for the panel columns k in [KB, KE) held in tensor memory from taddr on; `w` points at
// the window slot of block b-1.  Chunks and their window pairs are double-buffered: chunk c+1 is in flight while c is multiplied.
template <typename T, int NX, int R, int NCH, int KB, int KE>
__device__ __forceinline__ void tm_mv_part(uint32_t taddr, const T *__restrict__ w, T (&acc)[R][NCH])
{
    constexpr int V = TmWin<T, NX>::V;                    // elements per window load (4 floats or 2 doubles; 2 floats for nx = 2 mod 4)
    static_assert(NX % V == 0 && KB % V == 0, "window vectors never straddle a block");
    using Vec = typename TmWin<T, NX>::Vec;
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int W32 = (int)sizeof(T) / 4, CPK = R * W32, KPC = 16 / CPK, NK = KE - KB, NFULL = NK / KPC, REM = NK % KPC, NCHUNK = NFULL + (REM ? 1 : 0);
    static_assert(KPC % V == 0 && REM % V == 0, "whole window vectors per chunk");
    if constexpr (NCHUNK > 0) {
        uint32_t u[2][16];
        Vec x[2][KPC / V];
        auto issue = [&](int c, uint32_t (&dst)[16], Vec (&xd)[KPC / V]) {
            const int nk = (c < NFULL) ? KPC : REM;
            if (nk == KPC) tmem_ld16(taddr + 16 * c, dst);
            else tmem_ld8(taddr + 16 * c, dst);
#pragma unroll
            for (int h = 0; h < KPC / V; h++) {
                const int k = KB + c * KPC + V * h;
                if (V * h < nk) xd[h] = *reinterpret_cast<const Vec *>(w + (k / NX) * NXP + (k % NX));
            }
        };
        issue(0, u[0], x[0]);
#pragma unroll
        for (int c = 0; c < NCHUNK; c++) {
            const int nk = (c < NFULL) ? KPC : REM;
            if (nk == KPC) tmem_wait_ld(u[c & 1]);
            else tmem_wait_ld8(u[c & 1]);
            if (c + 1 < NCHUNK) issue(c + 1, u[(c + 1) & 1], x[(c + 1) & 1]);
#pragma unroll
            for (int kk = 0; kk < KPC; kk++) {
                if (kk < nk) {
                    const int k = KB + c * KPC + kk;
                    const T xv = tm_vec_get(x[c & 1][kk / V], kk % V);
#pragma unroll
                    for (int row = 0; row < R; row++) {
                        T m;
                        tm_unpack(&u[c & 1][(kk * R + row) * W32], m);
                        acc[row][k % NCH] = fma_t(m, xv, acc[row][k % NCH]);
                    }
                }
            }
        }
    }
}

// the same for the panel columns k in [0, KE) of a matrix whose rows are in registers
template <typename T, int NX, int R, int NCH, int KE>
__device__ __forceinline__ void tm_reg_part(const T (&M)[R][3 * NX], const T *__restrict__ w, T (&acc)[R][NCH])
{
    constexpr int V = TmWin<T, NX>::V;
    static_assert(NX % V == 0 && KE % V == 0, "window vectors never straddle a block");
    using Vec = typename TmWin<T, NX>::Vec;
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int GRP = 8;                                  // panel columns whose window vectors are loaded together
#pragma unroll
    for (int g = 0; g < KE; g += GRP) {
        Vec x[GRP / V];
#pragma unroll
        for (int h = 0; h < GRP / V; h++) {
            const int k = g + V * h;
            if (k < KE) x[h] = *reinterpret_cast<const Vec *>(w + (k / NX) * NXP + (k % NX));
        }
        asm volatile("" ::: "memory");
#pragma unroll
        for (int kk = 0; kk < GRP; kk++) {
            const int k = g + kk;
            if (k < KE) {
                const T xv = tm_vec_get(x[kk / V], kk % V);
#pragma unroll
                for (int row = 0; row < R; row++) acc[row][k % NCH] = fma_t(M[row][k], xv, acc[row][k % NCH]);
            }
        }
    }
}

#ifdef GBDPCG_TRACE
// phase timeline (scripts/ubench/trace_cluster.cu): clock64() of thread 0 of every CTA at the phase boundaries
__device__ long long g_trace[16][64][8];
#define FTRACE(m) do { if (t == 0 && it < 64 && blockIdx.x < 16) g_trace[blockIdx.x][it][m] = clock64(); } while (0)
// fixed costs of a solve: entry, init done, matrices arrived, rows in registers, loop entered, loop left, outputs written
__device__ long long g_trace_fix[16][8];
#define XTRACE(m) do { if (threadIdx.x == 0 && blockIdx.x < 16) g_trace_fix[blockIdx.x][m] = clock64(); } while (0)
#else
#define FTRACE(m) do { } while (0)
#define XTRACE(m) do { } while (0)
#endif


// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <typename T, int NX, int R, int PMODE, bool FUSED, bool CLUSTERED, int MAXT, int SK = 0>
__global__ void __launch_bounds__(MAXT, (PMODE == P_TMEM && MAXT <= 128) ? 2 : 1) pcg_fused_kernel(const KernelArgs<T> a)
{
    constexpr bool TM = (PMODE == P_TMEM);                                      // Phi^-1 rows (and SK columns of S) in tensor memory
    static_assert(!TM || (FUSED && MAXT <= 256 && (sizeof(T) == 8 ? R == 2 : (R == 4 && NX % 2 == 0))),
                  "P_TMEM: single-reduction recurrence, at most two warps per lane quarter; fp64 with two rows per thread, fp32 with four (even nx)");
    // Rows per thread need not divide the state size in the P_TMEM form: the last thread of a block-row then owns padding rows (nx 14, R 4:
    // rows 12, 13 and two that do not exist) -- zero matrix rows, zero vector entries, never loaded from or stored to global memory, never
    // pushed to a neighbour.  The window slot of a block is wide enough for them (NXP >= R * TPB).
    static_assert(NX % R == 0 || TM, "rows per thread must divide the state size");
    static_assert(Geo<T>::nxp(NX) >= R * ((NX + R - 1) / R), "the padding rows of the last thread must fit the window slot");
    static_assert(TM || SK == 0, "SK belongs to P_TMEM");
    static_assert(!TM || ((MAXT + 127) / 128) * TmGeo<T, NX, R, TM ? SK : 0>::CPT <= 512, "the warps of a lane quarter share its 512 tensor-memory columns");
    using TG = TmGeo<T, NX, R, TM ? SK : 0>;
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int BSZ = 3 * NX * NX;
    constexpr int TPB = (NX + R - 1) / R;                                       // threads per block-row
    constexpr bool PADR = (NX % R) != 0;                                        // the last thread of a block-row has padding rows
    constexpr int NWP = round_up(round_up(MAXT, 32) / 32, 2);                   // warp-partial pairs summed per reduction
    namespace cg = cooperative_groups;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = (int)a.N, bpc = (int)a.bpc;
    const bool use_tma = a.use_tma != 0;                                        // host decided: sources 16-byte aligned
    constexpr int nstage = (PMODE == P_REG) ? 2 : 1;                            // staging holds S, and Phi^-1 beside it when both go to registers
    const FusedPlan plan = make_fused_plan<T>(NX, NXP, bpc, PMODE, nstage);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *cred = reinterpret_cast<T *>(smem_raw + plan.cred_off);
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *ghost_r = reinterpret_cast<T *>(smem_raw + plan.ghost_off), *ghost_s = ghost_r + 2 * NXP, *ghost_w = ghost_r + 4 * NXP;
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);
    T *psm = reinterpret_cast<T *>(smem_raw + plan.psm_off);
    const uint32_t mbar = smem_u32(smem_raw + plan.mbar_off), xbar = smem_u32(smem_raw + plan.xbar_off);
    (void)xbar;
    const uint32_t mat_bytes_pad = (uint32_t)round_up(bpc * BSZ * (int)sizeof(T), 16);
    T *stageS = stage;
    T *stageP = (PMODE == P_SMEM) ? psm : reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(stage) + mat_bytes_pad);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    int rank = 0, csize = 1;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        rank = (int)cl.block_rank();
        csize = (int)cl.num_blocks();
    }
    const uint32_t cluster_id = blockIdx.x / (uint32_t)csize;
    const uint32_t n_clusters = gridDim.x / (uint32_t)csize;

    const int lb = t / TPB, ti = t - lb * TPB;            // local block-row, thread inside it
    const int r0 = ti * R;                                // first row (inside the block) of this thread
    const int b0 = rank * bpc;                            // first block-row of this CTA
    const int nb = max(0, min(bpc, N - b0));              // block-rows this CTA really owns
    const bool active = lb < nb;
    const int b = b0 + lb;
    const int last_rank = (N - 1) / bpc;

    auto sync_all = [&]() {
        if constexpr (CLUSTERED) cg::this_cluster().sync();
        else __syncthreads();
    };

    // one-time init
    for (int i = t; i < (bpc + 2) * NXP; i += blockDim.x) { winA[i] = (T)0; winB[i] = (T)0; }
    for (int i = t; i < 128; i += blockDim.x) red[i] = (T)0;
    for (int i = t; i < 64; i += blockDim.x) cred[i] = (T)0;
    if (t == 0 && use_tma) { mbar_init(mbar, 1); fence_mbar_init(); }
    // push mode: halo barriers [parity] at xbar, +8, reduction barriers [parity] at xbar+16, +24; one arrival (the owner's
    // expect_tx) per phase.  Phase bits live in registers and run on across the systems of a batch.
    uint32_t hb_phase = 0, rb_phase = 0;                   // bit `parity` = phase to wait for next
    if constexpr (CLUSTERED && FUSED && (FUSED_GHOST != 0) && (FUSED_PUSH != 0)) {
        if (t == 0) {
#pragma unroll
            for (int i = 0; i < 4; i++) mbar_init(xbar + 8 * i, 1);
            fence_mbar_init();
        }
    }

    // P_TMEM: warp 0 allocates the CTA's tensor-memory columns (a power of two that holds the warps of a lane quarter side by side)
    uint32_t *tbase_s = reinterpret_cast<uint32_t *>(smem_raw + plan.tbase_off);
    (void)tbase_s;
    if constexpr (TM) {
        if (warp == 0) {
            constexpr uint32_t ncols = (uint32_t)tm_alloc_cols(TG::CPT, MAXT);         // from MAXT, not blockDim: 256 columns for the half-size CTAs
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tbase_s)), "r"(ncols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }

    // window B sits at the same offset from window A in every CTA, so one pair of remote pointers serves both
    const int wdelta = (int)((plan.winB_off - plan.winA_off) / sizeof(T));
    T *winA_prev = nullptr, *winA_next = nullptr, *ghost_w_prev = nullptr, *ghost_w_next = nullptr;
    if constexpr (CLUSTERED) {
        cg::cluster_group cl = cg::this_cluster();
        if (rank > 0) { winA_prev = cl.map_shared_rank(winA, rank - 1); ghost_w_prev = cl.map_shared_rank(ghost_w, rank - 1); }
        if (rank + 1 < csize) { winA_next = cl.map_shared_rank(winA, rank + 1); ghost_w_next = cl.map_shared_rank(ghost_w, rank + 1); }
    }
    sync_all();
    uint32_t tbase = 0, tmine = 0;
    (void)tbase; (void)tmine;
    if constexpr (TM) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tbase = *tbase_s;
        tmine = tbase + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * TG::CPT);
    }

    const int slot = (lb + 1) * NXP + r0;                  // where this thread publishes its R vector elements
    const int wofs = (active ? lb : 0) * NXP;              // window [x_{b-1} | x_b | x_{b+1}] of this thread's block
    const bool send_prev = CLUSTERED && active && lb == 0 && rank > 0;
    const bool send_next = CLUSTERED && active && lb == nb - 1 && rank < last_rank;

    auto put = [&](T *dst, const T (&x)[R]) {
        if constexpr (R == 4 && sizeof(T) == 4) *reinterpret_cast<float4 *>(dst) = make_float4(x[0], x[1], x[2], x[3]);
        else if constexpr (R == 2 && sizeof(T) == 4) *reinterpret_cast<float2 *>(dst) = make_float2(x[0], x[1]);
        else if constexpr (R == 2 && sizeof(T) == 8) *reinterpret_cast<double2 *>(dst) = make_double2(x[0], x[1]);
        else {
#pragma unroll
            for (int i = 0; i < R; i++) dst[i] = x[i];
        }
    };
    auto publish = [&](const bool toB, const T (&x)[R], const bool remote = true) {
        const int wo = toB ? wdelta : 0;
        if (active) {
            put(winA + wo + slot, x);
            if constexpr (CLUSTERED) {
                if (remote && send_prev) put(winA_prev + wo + (bpc + 1) * NXP + r0, x);   // right halo of the previous CTA (always full)
                if (remote && send_next) put(winA_next + wo + r0, x);                     // left halo of the next CTA
            }
        }
    };
    // ghost copies (FUSED_GHOST): which halo sides of this CTA have a neighbour behind them
    const bool has_left = CLUSTERED && rank > 0 && nb > 0, has_right = CLUSTERED && rank < last_rank;
    (void)has_left; (void)has_right;

    // all-thread reduction of a (value0, value1) pair: bit-identical result in every thread of every CTA.
    // Consecutive reductions alternate between two slots, so a fast warp's next partial can never
    // overwrite what a slow warp is still summing.
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

    // single-value form for the two-reduction (textbook) order: half the shuffles and half the partials
    auto all_sum1 = [&](T x) -> T {
        T ws = warp_allsum(x);
        T *rd = red + red_slot * 64;
        if (lane == 0) rd[warp] = ws;
        __syncthreads();
        T s = tree_sum_smem<T, round_up(round_up(MAXT, 32) / 32, Geo<T>::V)>(rd);
        if constexpr (CLUSTERED) {
            T *cr = cred + red_slot * 32;
            if (t < csize) {
                T *dst = cg::this_cluster().map_shared_rank(cr, t);
                dst[rank] = s;
            }
            cg::this_cluster().sync();
            s = tree_sum_smem<T, 16>(cr);
        }
        red_slot ^= 1;
        return s;
    };

    // ---- matrix prefetch (TMA) ----
    auto issue_prefetch = [&](uint32_t prob) {
        // one thread: arm the barrier with the byte count, then launch the bulk copies of this CTA's block-rows
        const uint32_t bytes = (uint32_t)(nb * BSZ * (int)sizeof(T));
        const T *Sg = a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
        mbar_expect_tx(mbar, (PMODE != P_NONE) ? 2 * bytes : bytes);
        bulk_g2s(smem_u32(stageS), Sg, bytes, mbar);
        if constexpr (PMODE != P_NONE) {
            const T *Pg = a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ;
            bulk_g2s(smem_u32(stageP), Pg, bytes, mbar);
        }
    };
    // P_TMEM: the staging area holds ONE matrix at a time -- Phi^-1 (prefetched behind the previous system's iterations), then S
    auto issue_one = [&](const T *M, uint32_t prob) {
        const uint32_t bytes = (uint32_t)(nb * BSZ * (int)sizeof(T));
        mbar_expect_tx(mbar, bytes);
        bulk_g2s(smem_u32(stageS), M + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ, bytes, mbar);
    };
    (void)issue_one;
    uint32_t phase = 0;
    XTRACE(1);
    if (use_tma && t == 0 && nb > 0 && cluster_id < a.batch) {
        if constexpr (TM) issue_one(a.Pinv, cluster_id);
        else issue_prefetch(cluster_id);
    }

    for (uint32_t prob = cluster_id; prob < a.batch; prob += n_clusters) {
        const size_t vofs = (size_t)prob * a.vec_stride + (size_t)b * NX + r0;

        // vectors first (their latency overlaps the wait for the matrices)
        T lam[R], gam[R];
#pragma unroll
        for (int i = 0; i < R; i++) { lam[i] = (T)0; gam[i] = (T)0; }
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++)
                if (!PADR || r0 + i < NX) { lam[i] = a.lambda[vofs + i]; gam[i] = __ldg(a.gamma + vofs + i); }
        }

        // ---- matrices: HBM -> shared (TMA, or cooperative loads) -> registers ----
        T Sreg[R][3 * NX];
        T Preg[(PMODE == P_REG) ? R : 1][(PMODE == P_REG) ? 3 * NX : 1];
        (void)Preg;
        auto load_rows = [&](T (&M)[R][3 * NX], const T *img) {
            const T *row = img + (active ? lb * BSZ + r0 : 0);
#pragma unroll
            for (int k = 0; k < 3 * NX; k++) {
                if constexpr (R == 2 && sizeof(T) == 4 && FUSED_LOAD64) {
                    const float2 q = *reinterpret_cast<const float2 *>(row + k * NX);
                    M[0][k] = q.x; M[1][k] = q.y;
                } else if constexpr (R == 2 && sizeof(T) == 8) {
                    const double2 q = *reinterpret_cast<const double2 *>(row + k * NX);
                    M[0][k] = q.x; M[1][k] = q.y;
                } else if constexpr (R == 4 && sizeof(T) == 4 && NX % 4 == 0) {
                    const float4 q = *reinterpret_cast<const float4 *>(row + k * NX);
                    M[0][k] = q.x; M[1][k] = q.y; M[2][k] = q.z; M[3][k] = q.w;
                } else {
#pragma unroll
                    for (int i = 0; i < R; i++) M[i][k] = row[k * NX + i];
                }
            }
            // padding blocks of the layout (left of row 0, right of row N-1) are "never read" in the reference
            // (include/pcg.cuh:105-106): whatever they hold must not reach the arithmetic
            const bool kill_left = !active || b == 0, kill_right = !active || b == N - 1, kill_mid = !active;
#pragma unroll
            for (int k = 0; k < 3 * NX; k++) {
                const bool kill = (k < NX) ? kill_left : (k >= 2 * NX) ? kill_right : kill_mid;
#pragma unroll
                for (int i = 0; i < R; i++) M[i][k] = (kill || (PADR && r0 + i >= NX)) ? (T)0 : M[i][k];   // (padding rows: whatever was read is dropped)
            }
        };

        // Both paths end with the block-row images of S (and Phi^-1) in the staging area; the rows are then pulled
        // into registers from ONE place (two copies of that code overflow the register file).
        if constexpr (TM) {
            // Phi^-1: staging -> tensor memory (through the registers S will occupy), then S through the same staging area
            if (nb > 0) { mbar_wait(mbar, phase); phase ^= 1; }
            load_rows(Sreg, stageS);
            tm_store_rows<T, NX, R, 0, 3 * NX>(Sreg, tmine);
            __syncthreads();                               // every thread has read its Phi^-1 rows: staging is free
            if (t == 0 && nb > 0) { fence_proxy_async(); issue_one(a.S, prob); }
            if (nb > 0) { mbar_wait(mbar, phase); phase ^= 1; }
        } else if (use_tma) {
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
        XTRACE(2);
        load_rows(Sreg, stageS);
        if constexpr (PMODE == P_REG) load_rows(Preg, stageP);
        if constexpr (TM) {
            if constexpr (SK > 0) tm_store_rows<T, NX, R, TG::K0, 3 * NX>(Sreg, tmine + TG::PC);
            tmem_wait_st();
        }
        XTRACE(3);
        if constexpr (PMODE == P_SMEM) {
            // Phi^-1 stays in shared memory: clear its padding blocks in place (this CTA's first / last block-row)
            __syncthreads();
            if (b0 == 0) for (int i = t; i < NX * NX; i += blockDim.x) psm[i] = (T)0;
            if (nb > 0 && b0 + nb == N) for (int i = t; i < NX * NX; i += blockDim.x) psm[(nb - 1) * BSZ + 2 * NX * NX + i] = (T)0;
        }
        if (use_tma) {
            // staging may be overwritten by the next system's matrices once every thread holds its rows
            __syncthreads();
            const uint32_t next = prob + n_clusters;
            if constexpr (TM) {
                if (t == 0 && nb > 0 && next < a.batch) { fence_proxy_async(); issue_one(a.Pinv, next); }
            } else {
                if (t == 0 && nb > 0 && next < a.batch && PMODE != P_SMEM) { fence_proxy_async(); issue_prefetch(next); }
            }
        } else if constexpr (PMODE == P_SMEM) {
            __syncthreads();
        }
        const T *Prow = psm + (active ? lb * BSZ + r0 : 0);
        (void)Prow;

        // P_TMEM mat-vec: register part (S only), then the chunks in tensor memory; NCH chains per row, summed pairwise
        auto mv_tm = [&](const bool is_S, const T *wp, T (&y)[R]) {
            if constexpr (TM) {
                constexpr int NCH = FUSED_NCH_F64;
                T acc[R][NCH];
#pragma unroll
                for (int row = 0; row < R; row++)
#pragma unroll
                    for (int ch = 0; ch < NCH; ch++) acc[row][ch] = (T)0;
                if (is_S) {
                    tm_reg_part<T, NX, R, NCH, TG::K0>(Sreg, wp, acc);
                    tm_mv_part<T, NX, R, NCH, TG::K0, 3 * NX>(tmine + TG::PC, wp, acc);
                } else {
                    tm_mv_part<T, NX, R, NCH, 0, 3 * NX>(tmine, wp, acc);
                }
#pragma unroll
                for (int row = 0; row < R; row++) {
#pragma unroll
                    for (int n = NCH; n > 1; n = (n + 1) / 2) {
#pragma unroll
                        for (int i = 0; i < n / 2; i++) acc[row][i] = acc[row][i] + acc[row][i + (n + 1) / 2];
                    }
                    y[row] = acc[row][0];
                }
            }
        };
        (void)mv_tm;
        auto mv_S = [&](const T *win, T (&y)[R]) {
            if constexpr (TM) mv_tm(true, win + wofs, y);
            else block_row_mv<T, NX, R, false>(Sreg, nullptr, win + wofs, y);
            if (!active) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };
        auto mv_P = [&](const T *win, T (&y)[R]) {
            if constexpr (PMODE == P_REG) block_row_mv<T, NX, R, false>(Preg, nullptr, win + wofs, y);
            else if constexpr (PMODE == P_SMEM) block_row_mv<T, NX, R, true>(Sreg, Prow, win + wofs, y);
            else if constexpr (TM) mv_tm(false, win + wofs, y);
            if (!active) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };

        // ---- r = gamma - S*lambda                                               (pcg.cuh:118-126) ----
        // (the prologue uses whichever window the first loop publish does NOT touch, which saves a barrier)
        T res[R], rt[R], w[R], p[R];
        constexpr bool PRO_B = FUSED && PMODE != P_NONE;
        if constexpr (PRO_B) publish(true, lam);
        else publish(false, lam);
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
            // Cluster mode keeps GHOST copies of the neighbours' boundary block-rows of r (and of s): the neighbour sends its
            // boundary rows of w with the reduction's barrier, and r_g -= alpha s_g, s_g = w_g + beta s_g are repeated here
            // with the same alpha, beta and the same fused multiply-adds as on the owner, so the copies are bit-identical to
            // what an exchange would deliver -- and the publish of r needs only a block barrier.
            constexpr bool GHOST = CLUSTERED && (FUSED_GHOST != 0);
            // PUSH: what still crosses CTAs (halo rows of r~, boundary rows of w, reduction partials) is pushed with st.async
            // into the receiver's shared memory and counted on the receiver's mbarrier, so an exchange costs one DSMEM hop
            // instead of a release / barrier.cluster / acquire round trip.  Two barriers per purpose, alternating with the
            // iteration: a CTA can run at most one exchange ahead of its slowest partner (it needs that partner's partial
            // of the previous reduction), so a message can never land in a phase that is still being filled.
            constexpr bool PUSH = GHOST && (FUSED_PUSH != 0);
            constexpr bool R_IN_B = (PMODE == P_NONE);                 // the window that carries r
            constexpr int GE = (2 * NX + 31) / 32;                     // ghost elements per thread (blockDim.x >= 32)
            T *const rwin = R_IN_B ? winB : winA;
            T gr[GE], gs[GE];                                          // ghost r, s of the neighbours' boundary block-rows
            (void)gr; (void)gs;
            const uint32_t halo_bytes = (uint32_t)((has_left ? 1 : 0) + (has_right ? 1 : 0)) * NX * (uint32_t)sizeof(T);
            (void)halo_bytes;
            if constexpr (GHOST) {
                publish(R_IN_B, res);
                sync_all();
#pragma unroll
                for (int k = 0; k < GE; k++) {
                    const int e = t + k * (int)blockDim.x, sd = e / NX, c = e - sd * NX;
                    gr[k] = (e < 2 * NX) ? rwin[(sd ? bpc + 1 : 0) * NXP + c] : (T)0;
                    gs[k] = (T)0;
                }
            }
            auto push_rows = [&](T *local_dst, const uint32_t dst_rank, const uint32_t mb_local, const T (&x)[R]) {
                const uint32_t ra = mapa_u32(smem_u32(local_dst), dst_rank), rm = mapa_u32(mb_local, dst_rank);
                if constexpr (R == 2) st_async2(ra, x[0], x[1], rm);
                else {
#pragma unroll
                    for (int i = 0; i < R; i++)
                        if (!PADR || r0 + i < NX) st_async(ra + (uint32_t)(i * sizeof(T)), x[i], rm);   // (the receiver counts nx elements per side)
                }
            };
            for (;;) {
                FTRACE(0);
                const uint32_t par = it & 1u;
                if constexpr (PMODE != P_NONE) {
                    publish(false, res, !GHOST);
                    if constexpr (GHOST) __syncthreads();
                    else sync_all();
                    FTRACE(1);
                    mv_P(winA, rt);                        // r~ = Phi^-1 r
                    FTRACE(2);
                    if constexpr (PUSH) {
                        const uint32_t hb = xbar + 8 * par;
                        if (t == 0) mbar_expect_tx(hb, halo_bytes);
                        publish(true, rt, false);
                        if (send_prev) push_rows(winB + (bpc + 1) * NXP + r0, (uint32_t)(rank - 1), hb, rt);
                        if (send_next) push_rows(winB + r0, (uint32_t)(rank + 1), hb, rt);
                        __syncthreads();
                        mbar_wait_cluster(hb, (hb_phase >> par) & 1u);
                        hb_phase ^= 1u << par;
                    } else {
                        publish(true, rt);
                        sync_all();
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < R; i++) rt[i] = res[i];
                    publish(true, rt, !GHOST);
                    if constexpr (GHOST) __syncthreads();
                    else sync_all();
                }
                FTRACE(3);
                mv_S(winB, w);                             // w = S r~
                FTRACE(4);
                const uint32_t rb = xbar + 16 + 8 * par;
                (void)rb;
                if constexpr (PUSH) {
                    if (t == 0) mbar_expect_tx(rb, (uint32_t)csize * 2u * (uint32_t)sizeof(T) + halo_bytes);
                    if (send_prev) push_rows(ghost_w + (2 * par + 1) * NXP + r0, (uint32_t)(rank - 1), rb, w);
                    if (send_next) push_rows(ghost_w + 2 * par * NXP + r0, (uint32_t)(rank + 1), rb, w);
                } else if constexpr (GHOST) {
                    // boundary rows of w travel with the reduction's cluster barrier (slots alternate with the iteration)
                    if (send_prev) put(ghost_w_prev + (2 * par + 1) * NXP + r0, w);
                    if (send_next) put(ghost_w_next + 2 * par * NXP + r0, w);
                }
                T g = (T)0, d = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) { g += res[i] * rt[i]; d += w[i] * rt[i]; }
                Pair<T> gd;
                if constexpr (PUSH) {
                    // same arithmetic as all_sum(); the CTA partials are pushed to every CTA (itself included) instead of
                    // stored and fenced by a cluster barrier
                    Pair<T> pp(g, d);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) pp = pp + shfl_xor_pair<T>(pp, o);
                    T *rd = red + red_slot * 64;
                    if (lane == 0) {
                        if constexpr (sizeof(T) == 4) *reinterpret_cast<float2 *>(rd + 2 * warp) = make_float2(pp.x(), pp.y());
                        else *reinterpret_cast<double2 *>(rd + 2 * warp) = make_double2(pp.x(), pp.y());
                    }
                    __syncthreads();
                    const Pair<T> cs = pair_tree_sum<T, NWP>(rd);
                    T *cr = cred + red_slot * 32;
                    if (t < csize)
                        st_async2(mapa_u32(smem_u32(cr + 2 * rank), (uint32_t)t), cs.x(), cs.y(), mapa_u32(rb, (uint32_t)t));
                    mbar_wait_cluster(rb, (rb_phase >> par) & 1u);
                    rb_phase ^= 1u << par;
                    gd = pair_tree_sum<T, 16>(cr);
                    red_slot ^= 1;
                } else {
                    gd = all_sum(g, d);                    // gamma = r.r~ , delta = w.r~
                }
                FTRACE(5);
                const T gamma = gd.x(), delta = gd.y();
                // ghost rows of w: loaded now, needed only after alpha is known
                T gw[GE];
                (void)gw;
                if constexpr (GHOST) {
#pragma unroll
                    for (int k = 0; k < GE; k++) {
                        const int e = t + k * (int)blockDim.x, sd = e / NX, c = e - sd * NX;
                        gw[k] = (e < 2 * NX) ? ghost_w[(2 * par + sd) * NXP + c] : (T)0;
                    }
                }
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
                    lam[i] += alpha * p[i];
                    if constexpr (GHOST) {                 // the very operations the ghost copies repeat below
                        s[i] = fma_t(beta, s[i], w[i]);
                        res[i] = fma_t(-alpha, s[i], res[i]);
                    } else {
                        s[i] = w[i] + beta * s[i];
                        res[i] -= alpha * s[i];
                    }
                }
                if constexpr (GHOST) {
#pragma unroll
                    for (int k = 0; k < GE; k++) {
                        const int e = t + k * (int)blockDim.x, sd = e / NX, c = e - sd * NX;
                        if (e < 2 * NX && (sd ? has_right : has_left)) {
                            gs[k] = fma_t(beta, gs[k], gw[k]);
                            gr[k] = fma_t(-alpha, gs[k], gr[k]);
                            rwin[(sd ? bpc + 1 : 0) * NXP + c] = gr[k];
                        }
                    }
                }
                rcp_gamma = (T)1 / gamma;
                rcp_alpha = (T)1 / alpha;
                FTRACE(6);
                ++it;
            }
        } else {
            // ---------------- textbook order (reference include/pcg.cuh:131-208) ----------------
            if constexpr (PMODE != P_NONE) {
                publish(true, res);
                sync_all();
                mv_P(winB, rt);
            } else {
#pragma unroll
                for (int i = 0; i < R; i++) rt[i] = res[i];
            }
            T e0 = (T)0;
#pragma unroll
            for (int i = 0; i < R; i++) { p[i] = rt[i]; e0 += res[i] * rt[i]; }
            T eta = all_sum1(e0);
            for (; it < a.max_iter; ++it) {
                publish(false, p);
                sync_all();
                mv_S(winA, w);                             // upsilon = S p
                T v0 = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) v0 += p[i] * w[i];
                const T v = all_sum1(v0);
                const T alpha = eta / v;
#pragma unroll
                for (int i = 0; i < R; i++) { lam[i] += alpha * p[i]; res[i] -= alpha * w[i]; }
                if constexpr (PMODE != P_NONE) {
                    publish(true, res);
                    sync_all();
                    mv_P(winB, rt);                        // r~ = Phi^-1 r
                } else {
#pragma unroll
                    for (int i = 0; i < R; i++) rt[i] = res[i];
                }
                T e1 = (T)0;
#pragma unroll
                for (int i = 0; i < R; i++) e1 += res[i] * rt[i];
                const T eta_new = all_sum1(e1);
                if (abs_t(eta_new) < a.tol) { ++it; hit_max = false; break; }
                const T beta = eta_new / eta;
                eta = eta_new;
#pragma unroll
                for (int i = 0; i < R; i++) p[i] = rt[i] + beta * p[i];
            }
        }

        XTRACE(5);
        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++)
                if (!PADR || r0 + i < NX) a.lambda[vofs + i] = lam[i];
        }
        if (rank == 0 && t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        // No barrier here: every read of this system's windows and staging happened before the barrier inside
        // the last reduction, and the partial-sum slots alternate.
    }
    if constexpr (CLUSTERED) cg::this_cluster().sync();   // no CTA may exit while a neighbour can still store into it
    if constexpr (TM) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (warp == 0) {
            constexpr uint32_t ncols = (uint32_t)tm_alloc_cols(TG::CPT, MAXT);
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(ncols) : "memory");
        }
    }
    XTRACE(6);
}

}  // namespace gbdpcg
