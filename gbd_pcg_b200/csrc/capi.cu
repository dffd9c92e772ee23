// capi.cu -- host side of libgbdpcg.so: launch planning, workspaces and the extern "C" entry points
// declared in include/gbdpcg.h.  Replaces the host templates of the reference
// (include/interface.cuh:24-144) and its occupancy helper (include/pcg.cuh:13-49).
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "gbdpcg.h"
#include "kernel_table.h"
#include "pcg_fused.cuh"
#include "pcg_tmem4.cuh"
#include "pcg_generic.cuh"
#include "pcg_grid.cuh"
#include "pcg_rows.cuh"
#include "pinv_build.h"
#include "csr_pack.cuh"
#include "schur_form.cuh"
#include "sym_fill.cuh"

namespace gbdpcg {

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string tl_error;
static thread_local int tl_cuda_error = 0;
static std::atomic<uint64_t> g_launches{0};

static int fail(int status, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    tl_error = buf;
    return status;
}
static int fail_cuda(cudaError_t e, const char *what, int line)
{
    tl_cuda_error = (int)e;
    return fail(GBDPCG_ERR_CUDA, "CUDA error %d (%s) in %s at capi.cu:%d", (int)e, cudaGetErrorString(e), what, line);
}
#define CK(call)                                                         \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) return fail_cuda(e__, #call, __LINE__);  \
    } while (0)

// ------------------------------------------------------------------------------------------------
// launch planning
// ------------------------------------------------------------------------------------------------
static constexpr size_t kMaxDynSmem = 227 * 1024;  // sm_100: 227 KB opt-in dynamic shared memory per CTA

struct Plan {
    bool generic = false;
    bool fused_kernel = false;        // second-generation kernel (pcg_fused.cuh)
    bool tmem_kernel = false;         // tensor-memory batched kernel (pcg_tmem.cuh / pcg_tmem4.cuh)
    bool grid_one_barrier = false;    // ... its one-barrier form (pcg_grid1_kernel: other scratch layout, 2 TPB ghost-row threads)
    bool grid_kernel = false;         // cooperative-grid kernel for systems beyond one cluster (pcg_grid.cuh); `cluster` = CTAs of the grid
    int tmem_per_sm = 0;              // CTAs of it that share an SM (tensor-memory columns, registers)
    int rpt = 1, fused = 0;           // rows per thread, single-reduction recurrence
    const RowsVariant *v = nullptr;
    int cluster = 1, bpc = 0, threads = 0, nbuf = 1, pmode = P_NONE;
    int base_cluster = 1;             // cluster size the planner fitted before latency mode promoted it (pcg_fused)
    int ctas_per_sm = 1;              // P_TMEM form of pcg_fused: 2 for the CTAs of at most 128 threads (SMs per system = cluster / ctas_per_sm)
    size_t smem = 0;
    const void *fn = nullptr;
    const char *name = "";
};

static void all_variants(bool f64, int nx, const RowsVariant **tab, int *cnt)
{
    *tab = nullptr;
    *cnt = 0;
    switch (nx) {
#define CASE(NX) \
    case NX: *tab = f64 ? rows_variants_f64_##NX(cnt) : rows_variants_f32_##NX(cnt); break;
        GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
        default: break;
    }
}

static void fused_variants(bool f64, int nx, int fused, const FusedVariant **tab, int *cnt)
{
    *tab = nullptr;
    *cnt = 0;
    switch (nx) {
#define CASE(NX)                                                                                                       \
    case NX:                                                                                                           \
        *tab = f64 ? (fused ? fused_variants_f64_##NX##_1(cnt) : fused_variants_f64_##NX##_0(cnt))                     \
                   : (fused ? fused_variants_f32_##NX##_1(cnt) : fused_variants_f32_##NX##_0(cnt));                    \
        break;
        GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
#define CASE(NX)                                                                                                       \
    case NX:                                                                                                           \
        if (fused) *tab = f64 ? fused_variants_f64_##NX##_1(cnt) : fused_variants_f32_##NX##_1(cnt);                   \
        break;
        GBDPCG_FOR_EACH_NX_EXTRA(CASE)
#undef CASE
        default: break;
    }
}

// SMs of the current device (148 when there is none to ask: planning also works on a machine without a GPU)
static int device_sms()
{
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
        cudaGetLastError();
        return 148;
    }
    return sms;
}

// Second-generation kernel: Phi^-1 in registers when possible; the smallest cluster that fits -- or, when the whole batch
// stays resident in one wave anyway, a larger one (latency mode, below).
template <typename T>
static bool plan_fused(uint32_t nx, uint32_t N, bool has_p, const gbdpcg_config &cfg, uint32_t batch, Plan *out)
{
    constexpr bool f64 = sizeof(T) == 8;
    int fused = 1;                                             // single-reduction recurrence unless asked otherwise (parity measured per dtype)
    if (cfg.flags & GBDPCG_FLAG_TEXTBOOK) fused = 0;
    if (cfg.flags & GBDPCG_FLAG_FUSED) fused = 1;
    const FusedVariant *tab;
    int cnt;
    fused_variants(f64, (int)nx, fused, &tab, &cnt);
    if (cnt == 0) return false;
    // Placement preference first, cluster size second: Phi^-1 in registers on a larger cluster beats Phi^-1 in shared
    // memory on a smaller one (C4: 64 us on 16 CTAs vs 152 us on 8), so every cluster size is tried for P_REG before
    // P_SMEM is considered at all.
    static const int cl_opts[] = {1, 2, 4, 8, 16};
    int pm_opts[2] = {P_REG, P_SMEM};
    int n_pm = 2;
    if (!has_p) { pm_opts[0] = P_NONE; n_pm = 1; }
    else if (cfg.flags & GBDPCG_FLAG_PINV_SMEM) { pm_opts[0] = P_SMEM; n_pm = 1; }
    else if (cfg.flags & GBDPCG_FLAG_PINV_REG) { n_pm = 1; }
    for (int ip = 0; ip < n_pm; ip++) {
        for (int cl : cl_opts) {
            if (cfg.cluster && (int)cfg.cluster != cl) continue;
            const int bpc = ((int)N + cl - 1) / cl;
            for (int i = 0; i < cnt; i++) {
                const FusedVariant *v = &tab[i];
                const int threads = round_up(bpc * ((int)nx / v->rpt), 32);
                if (v->clustered != (cl > 1) || v->pmode != pm_opts[ip] || threads > v->maxt) continue;
                const size_t smem = make_fused_plan<T>((int)nx, v->nxp, bpc, v->pmode, v->pmode == P_REG ? 2 : 1).total;
                if (smem > kMaxDynSmem) continue;
                out->generic = false; out->fused_kernel = true; out->v = nullptr; out->cluster = cl; out->bpc = bpc;
                out->threads = threads; out->nbuf = (v->pmode == P_REG) ? 2 : 1; out->pmode = v->pmode; out->smem = smem;
                out->fn = v->fn; out->name = v->name; out->rpt = v->rpt; out->fused = v->fused; out->base_cluster = cl;
                // Latency mode.  A system that needs a cluster anyway solves faster on a LARGER one -- fewer rows per CTA,
                // and with the push-mode exchanges a CTA more costs nothing (C3: 63.5 us on 4 CTAs, 55.3 on 16; nx=24,N=64:
                // 88 -> 55; profiles/clusters_r1e.txt) -- as long as every system of the batch still gets its own cluster in
                // one wave and a CTA keeps at least two warps.  One CTA with Phi^-1 in registers stays one CTA (C2: 26.7 us
                // against 37), and large batches keep the smallest cluster: fewest SMs per system.
                if (cl > 1 && !cfg.cluster && v->fused) {
                    const unsigned long long sms = (unsigned long long)device_sms();
                    for (int big = 16; big > cl; big >>= 1) {
                        if ((unsigned long long)batch * big > sms) continue;
                        const int bpc2 = ((int)N + big - 1) / big;
                        const int threads2 = round_up(bpc2 * ((int)nx / v->rpt), 32);
                        if (threads2 < 64 || threads2 > v->maxt) continue;
                        const size_t smem2 = make_fused_plan<T>((int)nx, v->nxp, bpc2, v->pmode, v->pmode == P_REG ? 2 : 1).total;
                        if (smem2 > kMaxDynSmem) continue;
                        out->cluster = big; out->bpc = bpc2; out->threads = threads2; out->smem = smem2;
                        break;
                    }
                }
                return true;
            }
        }
    }
    return false;
}

// fp64 throughput form of the fused kernel (pcg_fused.cuh, PMODE = P_TMEM): two rows per thread, the rows of Phi^-1 and the last
// columns of the rows of S in tensor memory -- a system needs half the SMs of the register-row plan (C3: a 2-CTA cluster of 224
// threads instead of 4 CTAs).  The smallest cluster whose CTAs fit 256 threads, the shared-memory limit (staging holds one matrix
// at a time) and 512 tensor-memory columns.
// (fp32, even state sizes 8..16: the same form with FOUR rows per thread -- same words of S per thread, same columns; nx 10 and 14: the
// last thread of a block-row owns two padding rows)
template <typename T>
static bool plan_fused_tmem64(uint32_t nx, uint32_t N, bool has_p, const gbdpcg_config &cfg, Plan *out)
{
    constexpr bool f64 = sizeof(T) == 8;
    constexpr int RPT = f64 ? 2 : 4;
    {
        if (!has_p) return false;
        if (cfg.flags & (GBDPCG_FLAG_TEXTBOOK | GBDPCG_FLAG_PINV_SMEM | GBDPCG_FLAG_PINV_REG | GBDPCG_FLAG_NO_TMA | GBDPCG_FLAG_FORCE_GENERIC |
                         GBDPCG_FLAG_LEGACY_ROWS))
            return false;
        const FusedVariant *tab = nullptr;
        int cnt = 0, pair = 0;
        int sk4[4] = {0, 0, 0, 0};
        if (const char *e = getenv("GBDPCG_TMEM64_VARIANT")) pair = atoi(e) ? 1 : 0;      // tuning knob (A/B runs)
        if constexpr (f64) {
            switch (nx) {
#define CASE(NX) case NX: tab = fused_tmem_variants_f64_##NX(&cnt); for (int i = 0; i < 4; i++) sk4[i] = fused_tmem_sk_f64_##NX(i); break;
                GBDPCG_FOR_EACH_NX_TMEM64(CASE)
#undef CASE
                default: break;
            }
        } else {
            switch (nx) {
#define CASE(NX) case NX: tab = fused_tmem_variants_f32_##NX(&cnt); for (int i = 0; i < 4; i++) sk4[i] = fused_tmem_sk_f32_##NX(i); break;
                GBDPCG_FOR_EACH_NX_TMEM32(CASE)
#undef CASE
                default: break;
            }
        }
        if (cnt < 8) return false;
        // Two shapes of CTA at the smallest cluster that fits: up to 256 threads with the SM to itself, or -- when the CTAs have at most 128
        // threads anyway -- the 128-thread kernel with a 256-column allocation, so that TWO CTAs of two different systems share an SM and
        // cover each other's reduction and exchange phases (nx 16, N 32: 1.28 -> 0.84 ms per 2048).  Halving the CTAs on purpose (twice the
        // cluster) does not pay: C3 2.09 against 1.94 ms.  GBDPCG_TMEM64_SMALL=0 keeps one CTA per SM (A/B runs).
        const char *es = getenv("GBDPCG_TMEM64_SMALL");
        const bool allow_small = es ? atoi(es) != 0 : true;
        static const int cl_opts[] = {1, 2, 4, 8, 16};
        for (int cl : cl_opts) {
            if (cfg.cluster && (int)cfg.cluster != cl) continue;
            const int bpc = ((int)N + cl - 1) / cl;
            const int threads = round_up(bpc * (((int)nx + RPT - 1) / RPT), 32);
            for (int small = (allow_small && threads <= 128) ? 1 : 0; small >= 0; small--) {
                const int cpt = 2 * 2 * (3 * (int)nx + sk4[pair + 2 * small]);                // tensor-memory columns per thread
                const FusedVariant *v = &tab[4 * small + 2 * pair + (cl > 1 ? 1 : 0)];
                if (threads > v->maxt || ((v->maxt + 127) / 128) * cpt > 512) continue;
                const size_t smem = make_fused_plan<T>((int)nx, v->nxp, bpc, P_TMEM, 1).total;
                if (smem > kMaxDynSmem) continue;
                if (small && (2 * (smem + 1024) > 233472 || cpt > 256)) continue;              // two per SM or it is pointless
                out->generic = false; out->fused_kernel = true; out->tmem_kernel = false; out->grid_kernel = false; out->v = nullptr;
                out->cluster = cl; out->base_cluster = cl; out->bpc = bpc; out->threads = threads; out->nbuf = 1; out->pmode = P_TMEM;
                out->smem = smem; out->fn = v->fn; out->name = v->name; out->rpt = RPT; out->fused = 1; out->ctas_per_sm = small ? 2 : 1;
                return true;
            }
        }
        return false;
    }
}

// Systems too large for one 16-CTA cluster: a cooperative grid of register-row CTAs, at most one per SM (pcg_grid.cuh).
// The reference runs any N whose N CTAs are co-resident (include/pcg.cuh:23-49); this covers N up to SMs x block-rows per CTA.
template <typename T>
static bool plan_grid(uint32_t nx, uint32_t N, bool has_p, const gbdpcg_config &cfg, Plan *out, bool allow_one_barrier = true)
{
    constexpr bool f64 = sizeof(T) == 8;
    if (cfg.flags & (GBDPCG_FLAG_TEXTBOOK)) return false;      // single-reduction recurrence only
    if (cfg.cluster) return false;
    const GridVariant *tab = nullptr;
    int cnt = 0;
    switch (nx) {
#define CASE(NX) case NX: tab = f64 ? grid_variants_f64_##NX(&cnt) : grid_variants_f32_##NX(&cnt); break;
        GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
        default: break;
    }
    if (cnt == 0) return false;
    const int sms = device_sms();
    int pm_opts[2] = {P_REG, P_SMEM};
    int n_pm = 2;
    if (!has_p) { pm_opts[0] = P_NONE; n_pm = 1; }
    else if (cfg.flags & GBDPCG_FLAG_PINV_SMEM) { pm_opts[0] = P_SMEM; n_pm = 1; }
    else if (cfg.flags & GBDPCG_FLAG_PINV_REG) { n_pm = 1; }
    for (int ip = 0; ip < n_pm; ip++)
        for (int i = 0; i < cnt; i++) {
            const GridVariant *v = &tab[i];
            if (v->pmode != pm_opts[ip]) continue;
            const int tpb = (int)nx / v->rpt;
            int bpc_max = v->maxt / tpb;
            while (bpc_max > 0 && make_grid_plan<T>((int)nx, v->nxp, bpc_max, v->pmode).total > kMaxDynSmem) bpc_max--;
            int bpc = ((int)N + sms - 1) / sms;                    // as many CTAs as there are SMs: the iteration is barrier-bound, not work-bound
            const int bpc_2warps = (64 + tpb - 1) / tpb;           // ... but a CTA keeps at least two warps
            if (bpc < bpc_2warps) bpc = bpc_2warps;
            if (bpc > bpc_max) continue;
            // One barrier per iteration (pcg_grid1_kernel) where every CTA owns at least two block-rows and the CTA has room for the
            // 2 x tpb ghost-row threads.  GBDPCG_GRID_TWO_BARRIERS=1: A/B knob.
            auto one_barrier_fits = [&](int bb) {
                const int g1 = ((int)N + bb - 1) / bb, th1 = round_up(bb * tpb + 2 * tpb, 32);
                return allow_one_barrier && bb >= 2 && g1 >= 2 && th1 <= v->maxt && th1 >= 4 * (int)nx &&
                       make_grid1_plan<T>((int)nx, v->nxp, bb, v->pmode).total <= kMaxDynSmem && !getenv("GBDPCG_GRID_TWO_BARRIERS");
            };
            // ... and with one barrier HALF as many CTAs of twice the size are faster still where they fit (the barrier's arrivals are
            // serialised on one L2 line: nx 14 N 1500 fp32 126 -> 118 us on 69 CTAs, fp64 N 1000 184 -> 171 on 72; a third as many: 128).
            // GBDPCG_GRID_BPC_SCALE=1 keeps one CTA per SM (A/B knob).
            int scale = 2;
            if (const char *e = getenv("GBDPCG_GRID_BPC_SCALE")) { const int sc = atoi(e); if (sc >= 1 && sc <= 8) scale = sc; }
            if (scale > 1 && ((int)N + bpc - 1) / bpc > sms / 2 && bpc * scale <= bpc_max && one_barrier_fits(bpc * scale)) bpc *= scale;
            const int G = ((int)N + bpc - 1) / bpc;
            out->generic = false; out->fused_kernel = false; out->tmem_kernel = false; out->grid_kernel = true; out->v = nullptr;
            out->cluster = G; out->bpc = bpc; out->threads = round_up(bpc * tpb, 32); out->nbuf = 1; out->pmode = v->pmode;
            out->smem = make_grid_plan<T>((int)nx, v->nxp, bpc, v->pmode).total; out->fn = v->fn; out->name = v->name; out->rpt = v->rpt;
            out->fused = 1;
            out->grid_one_barrier = false;
            if (one_barrier_fits(bpc)) {
                out->grid_one_barrier = true;
                out->threads = round_up(bpc * tpb + 2 * tpb, 32);
                out->smem = make_grid1_plan<T>((int)nx, v->nxp, bpc, v->pmode).total; out->fn = v->fn1; out->name = v->name1;
            }
            return true;
        }
    return false;
}

template <typename T>
static size_t rows_smem(const RowsVariant *v, int bpc, int nbuf)
{
    return make_smem_plan<T>(v->nx, v->nxp, v->sb, bpc, nbuf).total;
}

const void *tmem_kernel_f32(int nx, const char **name);                                  // pcg_tmem_inst.cu
const void *tmem4_kernel_f32(int nx, int variant, const char **name, int *alloc_cols);   // pcg_tmem_inst.cu

// Shared memory a tensor-memory CTA asks for: at least so much that the hardware cannot place more CTAs on an SM than
// tensor memory has columns for (a surplus CTA would sit in tcgen05.alloc until a co-resident persistent CTA has
// finished its whole share of the batch: silent load imbalance).
static size_t tmem_smem_request(size_t need, int per_sm)
{
    const size_t sm_total = 233472;                                    // 228 KB per SM, 1 KB reserved per CTA
    const size_t floor_excl = sm_total / (size_t)(per_sm + 1);         // (per_sm + 1) x (smem + 1024) must exceed the SM
    const size_t min_req = floor_excl > 1024 ? floor_excl - 1024 + 16 : 0;
    return need > min_req ? need : min_req;
}

// Tensor-memory batched kernels: fp32, Phi^-1 given, even nx <= 16, one CTA per system, default recurrence.
//   pcg_tmem4.cuh: four rows per thread, <= 128 threads per system, two CTAs per SM (four when a CTA has <= 2 warps and
//                  <= 128 columns).  Taken when the system fits.
//   pcg_tmem.cuh:  two rows per thread, <= 256 threads per system (GBDPCG_FLAG_TMEM_R2 forces it).
//   pcg_tmem4.cuh, wide form: four rows per thread, 129..256 threads per system (nx 14: 37 <= N <= 64), one CTA per SM.
// Worth it only when the batch fills every slot of the GPU.
static bool plan_tmem(uint32_t nx, uint32_t N, bool has_p, const gbdpcg_config &cfg, uint32_t batch, Plan *out)
{
    if (!has_p || (nx & 1u) || nx > 16 || N < 2 || cfg.cluster > 1) return false;
    if (cfg.flags & (GBDPCG_FLAG_TEXTBOOK | GBDPCG_FLAG_PINV_SMEM | GBDPCG_FLAG_NO_TMA)) return false;
    const char *name = nullptr;
    if (!(cfg.flags & GBDPCG_FLAG_TMEM_R2)) {
        const int threads4 = round_up((int)(N * ((nx + 3) / 4)), 32);
        int variant = threads4 <= 64 ? 10 : 0, cols = 0;                                        // <= 2 warps: the <= 128-column variant, four CTAs per SM
        if (const char *e = getenv("GBDPCG_TMEM4_VARIANT")) variant = atoi(e);                  // tuning knob (A/B runs)
        const void *fn4 = tmem4_kernel_f32((int)nx, variant, &name, &cols);
        if (fn4 && threads4 <= kTmem4MaxThreads) {
            int per_sm = 512 / cols;
            const int by_regs = 65536 / (256 * threads4);                                      // __launch_bounds__(128, 2): up to 255 registers
            if (by_regs < per_sm) per_sm = by_regs;
            const size_t need = make_tmem4_plan((int)nx, t4_nxp((int)nx), (int)N).total;
            // Solo form: a launch of at most one system per SM (single solves, small batches) gives S a staging image of its own, so both
            // matrices of a system are in flight at once and nothing of the next system arrives exposed (C2 single solve 22.6 us packed
            // form against 26.7 for the register-row kernel, profiles/single_tmem4_r2f.txt).
            const size_t solo = need + t4_stage_bytes((int)nx, (int)N);
            if (threads4 >= 96 && batch <= (uint32_t)device_sms() && solo <= kMaxDynSmem && !getenv("GBDPCG_TMEM4_NO_SOLO")) {
                out->generic = false; out->fused_kernel = false; out->tmem_kernel = true; out->tmem_per_sm = 1; out->v = nullptr;
                out->cluster = 1; out->bpc = (int)N; out->threads = threads4; out->nbuf = 2; out->pmode = P_REG;
                out->smem = tmem_smem_request(solo, 1); out->fn = fn4; out->name = name; out->rpt = 4; out->fused = 1;
                return true;
            }
            // worth it with at least 8 warps on an SM (nx=12,N=20 at 2 x 2 warps: 0.48 ms against 0.40 for the two-row kernel; profiles/tmem4_r2a.txt)
            const bool enough_warps = per_sm * threads4 >= 256 || (cfg.flags & GBDPCG_FLAG_TMEM) || getenv("GBDPCG_TMEM4_VARIANT");
            if (enough_warps && (size_t)per_sm * (need + 1024) <= 233472) {
                out->generic = false; out->fused_kernel = false; out->tmem_kernel = true; out->tmem_per_sm = per_sm; out->v = nullptr;
                out->cluster = 1; out->bpc = (int)N; out->threads = threads4; out->nbuf = 1; out->pmode = P_REG;
                out->smem = tmem_smem_request(need, per_sm); out->fn = fn4; out->name = name; out->rpt = 4; out->fused = 1;
                return true;
            }
        }
    }
    const void *fn = tmem_kernel_f32((int)nx, &name);
    if (!fn) return false;
    const int threads = round_up((int)(N * nx / 2), 32);
    if (threads > kTmemMaxThreads) {
        // beyond the two-row kernel: the WIDE four-row form, 8 warps (two per lane quarter), one CTA per SM (nx 14: N <= 64)
        const int threads4 = round_up((int)(N * ((nx + 3) / 4)), 32);
        int cols = 0;
        const void *fnw = (cfg.flags & GBDPCG_FLAG_TMEM_R2) ? nullptr : tmem4_kernel_f32((int)nx, 20, &name, &cols);
        const size_t need = make_tmem4_plan((int)nx, t4_nxp((int)nx), (int)N).total;
        if (!fnw || threads4 > kTmem4WideThreads || need > kMaxDynSmem) return false;
        out->generic = false; out->fused_kernel = false; out->tmem_kernel = true; out->tmem_per_sm = 1; out->v = nullptr; out->cluster = 1;
        out->bpc = (int)N; out->threads = threads4; out->nbuf = 1; out->pmode = P_REG; out->smem = tmem_smem_request(need, 1); out->fn = fnw;
        out->name = name; out->rpt = 4; out->fused = 1;
        return true;
    }
    const size_t smem = make_tmem_plan((int)nx, Geo<float>::nxp((int)nx), (int)N).total;
    const int per_sm = 512 / tmem_cols(threads);
    if ((size_t)per_sm * (smem + 1024) > 233472) return false;   // as many CTAs per SM as tensor memory allows, or it is pointless
    out->generic = false; out->fused_kernel = false; out->tmem_kernel = true; out->tmem_per_sm = per_sm; out->v = nullptr; out->cluster = 1;
    out->bpc = (int)N; out->threads = threads; out->nbuf = 1; out->pmode = P_REG; out->smem = tmem_smem_request(smem, per_sm); out->fn = fn;
    out->name = name; out->rpt = 2; out->fused = 1;
    return true;
}

static int prepare_fn(const void *fn, bool clustered, bool max_carveout = false);   // below: per-(device, kernel) attributes
static int max_resident_problems(const Plan &p, int *out);                         // below: occupancy of a plan on the current device

template <typename T>
static int make_plan(uint32_t nx, uint32_t N, bool has_p, const gbdpcg_config &cfg, uint32_t batch, Plan *out)
{
    constexpr bool f64 = sizeof(T) == 8;
    const RowsVariant *tab;
    int cnt;
    all_variants(f64, (int)nx, &tab, &cnt);
    if (cfg.flags & GBDPCG_FLAG_TMEM) {
        if constexpr (!f64) { if (!cfg.cluster && plan_tmem(nx, N, has_p, cfg, batch, out)) return GBDPCG_OK; }
        if (!(cfg.flags & GBDPCG_FLAG_TMEM_R2) && plan_fused_tmem64<T>(nx, N, has_p, cfg, out)) return GBDPCG_OK;   // (TMEM_R2 asks for the two-row kernel and nothing else)
        return fail(GBDPCG_ERR_UNSUPPORTED, "GBDPCG_FLAG_TMEM: Phi^-1 given and fp32, even nx <= 16, N*ceil(nx/4) <= 256 threads -- or fp64, even nx 8..16, "
                    "N*nx/2 <= 16 x 256 threads (nx=%u N=%u %s)", nx, N, f64 ? "f64" : "f32");
    }
    // Throughput mode: a batch that fills every tensor-memory slot of the GPU (two systems per SM above 128 threads per
    // system, four below) goes to the tensor-memory kernel -- same arithmetic bit for bit (C5: 1.18 -> 1.00 ms; nx=16,N=30:
    // 1.43 -> 1.00; nx=10,N=17: 0.42 -> 0.35; profiles/tmem_r1g.txt).
    if constexpr (!f64) {
        const uint32_t auto_off = GBDPCG_FLAG_NO_TMEM | GBDPCG_FLAG_FORCE_GENERIC | GBDPCG_FLAG_LEGACY_ROWS | GBDPCG_FLAG_PINV_REG | GBDPCG_FLAG_PINV_SMEM |
                                  GBDPCG_FLAG_TEXTBOOK | GBDPCG_FLAG_NO_TMA;
        // (two-row kernel: systems of <= 64 threads lose, the register-row kernel already fits four of them on an SM -- nx=8,N=16: 0.19 vs 0.21 ms)
        // Four rows per thread, systems of 3-4 warps: every batch size -- above #SMs systems two CTAs per SM, up to #SMs the solo form
        // (also the single solve).  Smaller systems and the two-row kernel: only when the batch fills every tensor-memory slot.
        Plan tp;
        if (!(cfg.flags & auto_off) && !cfg.cluster && has_p && plan_tmem(nx, N, has_p, cfg, batch, &tp)) {
            const bool wide = tp.rpt == 4 && tp.threads > kTmem4MaxThreads;
            const bool big4 = tp.rpt == 4 && tp.threads >= 96 && !wide;
            const bool fills = (unsigned long long)batch >= (unsigned long long)tp.tmem_per_sm * (unsigned long long)device_sms();
            // the wide form (one system per SM) replaces register-row CLUSTERS in throughput mode only: a batch that cannot be resident at once
            bool wide_wins = false;
            if (wide) {
                Plan rp;
                // (against a register-row plan of ONE CTA it is still 10-15 % ahead once every SM has a system: nx 8, N 80: 0.62 -> 0.55 ms)
                wide_wins = plan_fused<T>(nx, N, has_p, cfg, batch, &rp) &&
                            (rp.base_cluster > 1 ? (unsigned long long)batch * (unsigned long long)rp.base_cluster > (unsigned long long)device_sms()
                                                 : (unsigned long long)batch >= (unsigned long long)device_sms());
            }
            if (big4 || wide_wins || (!wide && fills && (tp.rpt == 4 || N * nx / 2 > 64))) {
                *out = tp;
                return GBDPCG_OK;
            }
        }
    }
    // Throughput mode of everything the kernels above do not take (fp64; fp32 systems beyond 256 four-row threads, nx a multiple of 4): a
    // batch whose register-row clusters cannot all be resident at once takes the tensor-memory form of the fused kernel when that needs
    // fewer SMs per system (C3 x 2048: 2-CTA clusters instead of 4-CTA ones, 3.61 -> 1.94 ms; fp32 C4 x 2048: 8 CTAs instead of 16,
    // 9.62 -> 5.78 ms; nx 8 x N 200 fp32: 2.23 -> 1.01 ms)
    {
        const uint32_t auto_off = GBDPCG_FLAG_NO_TMEM | GBDPCG_FLAG_FORCE_GENERIC | GBDPCG_FLAG_LEGACY_ROWS | GBDPCG_FLAG_PINV_REG | GBDPCG_FLAG_PINV_SMEM |
                                  GBDPCG_FLAG_TEXTBOOK | GBDPCG_FLAG_NO_TMA;
        Plan tp, rp;
        if (!(cfg.flags & auto_off) && !cfg.cluster && has_p && plan_fused_tmem64<T>(nx, N, has_p, cfg, &tp)) {
            const bool have_reg = plan_fused<T>(nx, N, has_p, cfg, batch, &rp);
            // Beyond what a 16-CTA cluster of register rows holds, a 16-CTA cluster of this form still does (nx 14 fp32: N <= 1024 instead of 752,
            // fp64 N <= 576 instead of 288): it goes before the cooperative grid -- nx 14 x N 800 fp32 single solve 148.8 -> 58.5 us.
            if (!have_reg) {
                *out = tp;
                return GBDPCG_OK;
            }
            const bool throughput = (unsigned long long)batch * (unsigned long long)(have_reg ? rp.base_cluster : 1) > (unsigned long long)device_sms();
            // (equal cluster sizes: still ahead where the register-row plan has to keep Phi^-1 in shared memory -- nx 16, N 32: 2.14 -> 1.29 ms)
            const int tp_sms2 = 2 * tp.cluster / tp.ctas_per_sm, rp_sms2 = 2 * rp.base_cluster;      // SMs per system, doubled
            // (fp32: only where the register rows need a cluster -- small fp32 systems already run several per SM on the register-row kernel)
            const bool in_domain = f64 || rp.base_cluster > 1;
            if (have_reg && throughput && in_domain && (tp_sms2 < rp_sms2 || (tp_sms2 == rp_sms2 && rp.pmode == P_SMEM))) {
                *out = tp;
                return GBDPCG_OK;
            }
            // Latency mode (every system of the batch gets its own cluster in one wave): the tensor-memory form on the LARGEST cluster whose CTAs
            // keep two warps -- two rows per thread halve the threads a CTA has to synchronise and the window traffic.  Single solves, p50 against
            // the register-row latency plan: C3 55.4 -> 49.4 us, nx 8 x N 200 57.4 -> 45.1, nx 14 x N 128 59.6 -> 53.2, nx 12 x N 100 51.2 -> 49.3;
            // nx 10 and 16 measured 2-4 % behind, they keep the register rows.  fp32 (four rows per thread): nx 8 and 16 gain 13-16 % (nx 8 x N 200
            // 38.9 -> 32.9 us, nx 8 x N 512 41.0 -> 34.8, nx 16 x N 100 45.1 -> 39.0, nx 16 x N 64 45.2 -> 38.9), nx 12 loses (N 100: 30.7 -> 34.8; C4 equal).
            const bool latency_ok = f64 ? (nx != 10 && nx != 16) : (nx == 8 || nx == 16);
            if (!throughput && latency_ok && !getenv("GBDPCG_TMEM64_NO_LATENCY")) {
                for (int cl = 16; cl >= 2; cl >>= 1) {
                    if ((unsigned long long)batch * (unsigned long long)cl > (unsigned long long)device_sms()) continue;
                    const int bpc = ((int)N + cl - 1) / cl;
                    if (round_up(bpc * (((int)nx + (f64 ? 1 : 3)) / (f64 ? 2 : 4)), 32) < 64 || (cl - 1) * bpc >= (int)N) continue;   // two warps per CTA, no empty CTA
                    gbdpcg_config c2 = cfg;
                    c2.cluster = (uint32_t)cl;
                    Plan lp;
                    int res = 0;
                    if (!plan_fused_tmem64<T>(nx, N, has_p, c2, &lp)) continue;
                    // the whole batch must be resident at once with this cluster, or the register-row plan (which steps down by itself) stays
                    if (prepare_fn(lp.fn, lp.cluster > 1, lp.ctas_per_sm > 1) != GBDPCG_OK || max_resident_problems(lp, &res) != GBDPCG_OK || (uint32_t)res < batch) {
                        cudaGetLastError();
                        break;
                    }
                    *out = lp;
                    return GBDPCG_OK;
                }
            }
        }
    }
    if (!(cfg.flags & (GBDPCG_FLAG_FORCE_GENERIC | GBDPCG_FLAG_LEGACY_ROWS)) && plan_fused<T>(nx, N, has_p, cfg, batch, out)) return GBDPCG_OK;
    // beyond one cluster: the cooperative-grid kernel, before the first-generation kernels get a say
    if (!(cfg.flags & (GBDPCG_FLAG_FORCE_GENERIC | GBDPCG_FLAG_LEGACY_ROWS)) && plan_grid<T>(nx, N, has_p, cfg, out)) return GBDPCG_OK;
    if (cnt > 0 && !(cfg.flags & GBDPCG_FLAG_FORCE_GENERIC)) {
        static const int cl_opts[] = {1, 2, 4, 8, 16};
        for (int cl : cl_opts) {
            if (cfg.cluster && (int)cfg.cluster != cl) continue;
            const int bpc = ((int)N + cl - 1) / cl;
            const int threads = round_up(bpc * (int)nx, 32);
            // Phi^-1 placement preference: registers, then shared memory
            int pm_opts[2] = {P_REG, P_SMEM};
            int n_pm = 2;
            if (!has_p) { pm_opts[0] = P_NONE; n_pm = 1; }
            else if (cfg.flags & GBDPCG_FLAG_PINV_SMEM) { pm_opts[0] = P_SMEM; n_pm = 1; }
            else if (cfg.flags & GBDPCG_FLAG_PINV_REG) { n_pm = 1; }
            for (int ip = 0; ip < n_pm; ip++) {
                for (int i = 0; i < cnt; i++) {
                    const RowsVariant *v = &tab[i];
                    if (v->clustered != (cl > 1) || v->pmode != pm_opts[ip] || threads > v->maxt) continue;
                    int nbuf = (v->pmode == P_NONE) ? 1 : 2;
                    size_t smem = rows_smem<T>(v, bpc, nbuf);
                    if (smem > kMaxDynSmem && nbuf == 2) { nbuf = 1; smem = rows_smem<T>(v, bpc, 1); }
                    if (smem > kMaxDynSmem) continue;
                    out->generic = false; out->v = v; out->cluster = cl; out->bpc = bpc; out->threads = threads;
                    out->nbuf = nbuf; out->pmode = v->pmode; out->smem = smem; out->fn = v->fn; out->name = v->name;
                    return GBDPCG_OK;
                }
            }
        }
        if (cfg.cluster || (cfg.flags & (GBDPCG_FLAG_PINV_SMEM | GBDPCG_FLAG_PINV_REG)))
            return fail(GBDPCG_ERR_UNSUPPORTED, "no register-row kernel fits nx=%u N=%u %s with the forced cluster=%u flags=0x%x",
                        nx, N, f64 ? "f64" : "f32", cfg.cluster, cfg.flags);
    }
    // runtime-nx fallback: one CTA, vectors in shared memory
    const size_t smem = generic_smem_bytes<T>(nx, N);
    if (smem > kMaxDynSmem)
        return fail(GBDPCG_ERR_UNSUPPORTED, "system nx=%u N=%u %s needs %zu B of shared memory in the fallback kernel (limit %zu)",
                    nx, N, f64 ? "f64" : "f32", smem, kMaxDynSmem);
    out->generic = true; out->v = nullptr; out->cluster = 1; out->bpc = (int)N;
    const int rows = (int)(nx * N);
    out->threads = rows >= 1024 ? 1024 : round_up(rows, 32);
    out->nbuf = 0; out->pmode = has_p ? 3 : 0; out->smem = smem;
    out->fn = has_p ? (const void *)pcg_generic_kernel<T, true> : (const void *)pcg_generic_kernel<T, false>;
    out->name = has_p ? (f64 ? "pcg_generic<f64,P>" : "pcg_generic<f32,P>") : (f64 ? "pcg_generic<f64,I>" : "pcg_generic<f32,I>");
    return GBDPCG_OK;
}

// per-(device, kernel) one-time attribute setup + occupancy
struct FnState { const void *fn; int device; int clustered; };
static std::mutex g_fn_mu;
static std::vector<FnState> g_fn_ready;
struct OccState { const void *fn; int device, cluster, threads; size_t smem; int resident; };
static std::vector<OccState> g_occ;

static int prepare_fn(const void *fn, bool clustered, bool max_carveout)
{
    int dev;
    CK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_fn_mu);
    for (auto &s : g_fn_ready)
        if (s.fn == fn && s.device == dev) return GBDPCG_OK;
    CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
    if (clustered) CK(cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    // half-size CTAs of the tensor-memory form: two of ~80 KB must be able to share an SM (the default carveout for that request holds one)
    if (max_carveout) CK(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    g_fn_ready.push_back({fn, dev, clustered});
    return GBDPCG_OK;
}

static int max_resident_problems(const Plan &p, int *out)
{
    int dev, sms;
    CK(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(g_fn_mu);
        for (auto &o : g_occ)
            if (o.fn == p.fn && o.device == dev && o.cluster == p.cluster && o.threads == p.threads && o.smem == p.smem) {
                *out = o.resident;
                return GBDPCG_OK;
            }
    }
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (p.tmem_kernel) {
        *out = p.tmem_per_sm * sms;                        // bounded by tensor-memory columns and registers (plan_tmem), enforced through the shared-memory request
        return GBDPCG_OK;
    }
    if (p.grid_kernel) {
        int per_sm = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, p.fn, p.threads, p.smem));
        *out = (per_sm * sms >= p.cluster) ? 1 : 0;       // the whole grid must be co-resident (cooperative launch); one system at a time
    } else if (p.cluster == 1) {
        int per_sm = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, p.fn, p.threads, p.smem));
        *out = per_sm * sms * p.ctas_per_sm;
    } else {
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3(p.cluster);
        lc.blockDim = dim3(p.threads);
        lc.dynamicSmemBytes = p.smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = p.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        lc.attrs = at; lc.numAttrs = 1;
        int ncl = 0;
        CK(cudaOccupancyMaxActiveClusters(&ncl, p.fn, &lc));
        *out = ncl * p.ctas_per_sm;
    }
    std::lock_guard<std::mutex> lk(g_fn_mu);
    g_occ.push_back({p.fn, dev, p.cluster, p.threads, p.smem, *out});
    return GBDPCG_OK;
}

// Stream-ordered scratch for the cooperative-grid kernel comes from a pool of this library's own, one per device, that keeps its
// memory between launches: the device's default pool releases everything at the next synchronisation point (release threshold 0),
// which made every launch pay ~0.75 ms of allocator work (profiles/grid_r2b.txt).
static int scratch_pool(cudaMemPool_t *out)
{
    static std::mutex mu;
    static std::vector<std::pair<int, cudaMemPool_t>> pools;
    int dev;
    CK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    for (auto &e : pools)
        if (e.first == dev) { *out = e.second; return GBDPCG_OK; }
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    cudaMemPool_t pool;
    CK(cudaMemPoolCreate(&pool, &props));
    unsigned long long keep = ~0ull;
    CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    pools.push_back({dev, pool});
    *out = pool;
    return GBDPCG_OK;
}

template <typename T>
static int launch(const Plan &p, const T *S, const T *P, const T *gamma, T *lambda, uint32_t nx, uint32_t N,
                  uint32_t batch, const gbdpcg_config &cfg, uint32_t *d_iters, uint8_t *d_exits, cudaStream_t stream)
{
    if (batch == 0) return GBDPCG_OK;
    int rc = prepare_fn(p.fn, p.cluster > 1 && !p.grid_kernel, p.ctas_per_sm > 1);
    if (rc) return rc;
    int resident = 0;
    rc = max_resident_problems(p, &resident);
    if (rc) return rc;
    if (p.fused_kernel && p.cluster > p.base_cluster && (resident < 1 || (uint32_t)resident < batch)) {
        // Latency mode promoted this plan to a larger (possibly non-portable 16-CTA) cluster on the strength of SM arithmetic alone.
        // If that many clusters cannot be co-resident after all (MIG slice, fewer SMs per GPC, another tenant on the GPCs), step
        // down towards the cluster size the planner had fitted first instead of failing or running a second pass.
        Plan q = p;
        while (q.cluster > p.base_cluster) {
            q.cluster >>= 1;
            q.bpc = ((int)N + q.cluster - 1) / q.cluster;
            q.threads = round_up(q.bpc * ((int)nx / q.rpt), 32);
            q.smem = make_fused_plan<T>((int)nx, Geo<T>::nxp((int)nx), q.bpc, q.pmode, q.pmode == P_REG ? 2 : 1).total;
            int r2 = 0;
            if (max_resident_problems(q, &r2) == GBDPCG_OK && r2 >= 1 && ((uint32_t)r2 >= batch || q.cluster == p.base_cluster))
                return launch<T>(q, S, P, gamma, lambda, nx, N, batch, cfg, d_iters, d_exits, stream);
        }
    }
    if (resident < 1)
        return fail(GBDPCG_ERR_UNSUPPORTED, "kernel %s cannot become resident (cluster=%d threads=%d smem=%zu)", p.name, p.cluster,
                    p.threads, p.smem);
    if (p.tmem_kernel) {
        // 128 or 256 of the SM's 512 tensor-memory columns per CTA: four or two CTAs per SM; bulk copies only
        {
            // the occupancy query answers 1 per SM for 2 x 80 KB unless the kernel asks for the full shared-memory carveout
            static std::mutex mu;
            static std::vector<std::pair<const void *, int>> done;
            int dev;
            CK(cudaGetDevice(&dev));
            std::lock_guard<std::mutex> lk(mu);
            bool seen = false;
            for (auto &d : done) seen = seen || (d.first == p.fn && d.second == dev);
            if (!seen) {
                CK(cudaFuncSetAttribute(p.fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
                done.push_back({p.fn, dev});
            }
        }
        if (((uintptr_t)S % 16) || ((uintptr_t)P % 16) || ((3ull * N * nx * nx * sizeof(T)) % 16)) {
            if (cfg.flags & GBDPCG_FLAG_TMEM) return fail(GBDPCG_ERR_UNSUPPORTED, "GBDPCG_FLAG_TMEM needs 16-byte aligned S and Pinv");
            // chosen by the planner: fall back to the register-row kernel, which can stage with ordinary loads
            gbdpcg_config c2 = cfg;
            c2.flags |= GBDPCG_FLAG_NO_TMEM;
            Plan p2;
            if ((rc = make_plan<T>(nx, N, P != nullptr, c2, batch, &p2))) return rc;
            return launch<T>(p2, S, P, gamma, lambda, nx, N, batch, c2, d_iters, d_exits, stream);
        }
    }
    if (p.grid_kernel) {
        // one cooperative grid solves the systems of the batch one after the other; what crosses CTAs goes through a small
        // stream-ordered scratch allocation (concurrent callers on other streams get their own)
        KernelArgs<T> a;
        a.S = S; a.Pinv = P; a.gamma = gamma; a.lambda = lambda; a.iters = d_iters; a.exits = d_exits;
        a.mat_stride = 3ull * N * nx * nx; a.vec_stride = (unsigned long long)N * nx;
        a.N = N; a.batch = batch; a.bpc = (uint32_t)p.bpc; a.max_iter = cfg.max_iter; a.two_buffers = 0; a.use_tma = 0;
        a.tol = (T)cfg.exit_tol;
        const int G = p.cluster;
        T *scratch = nullptr;
        cudaMemPool_t pool;
        if ((rc = scratch_pool(&pool))) return rc;
        const size_t selems = p.grid_one_barrier ? grid1_scratch_elems<T>((int)nx, G) : grid_scratch_elems<T>((int)nx, G);
        CK(cudaMallocFromPoolAsync((void **)&scratch, selems * sizeof(T), pool, stream));
        GridScratch<T> gs;
        gs.halo_r = scratch; gs.halo_rt = scratch + (size_t)G * 2 * nx; gs.halo_w = scratch + (size_t)G * 4 * nx;
        gs.partials = scratch + (size_t)G * 6 * nx;
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3((unsigned)G);
        lc.blockDim = dim3(p.threads);
        lc.dynamicSmemBytes = p.smem;
        lc.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeCooperative;
        at[0].val.cooperative = 1;
        lc.attrs = at; lc.numAttrs = 1;
        void *args[] = {&a, &gs};
        void *args1[] = {&a, &scratch};
        const cudaError_t e = cudaLaunchKernelExC(&lc, p.fn, p.grid_one_barrier ? args1 : args);
        cudaFreeAsync(scratch, stream);
        if (e != cudaSuccess) return fail_cuda(e, "cudaLaunchKernelExC (cooperative grid)", __LINE__);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return GBDPCG_OK;
    }
    const uint32_t n_groups = batch < (uint32_t)resident ? batch : (uint32_t)resident;

    KernelArgs<T> a;
    a.S = S; a.Pinv = P; a.gamma = gamma; a.lambda = lambda; a.iters = d_iters; a.exits = d_exits;
    a.mat_stride = 3ull * N * nx * nx; a.vec_stride = (unsigned long long)N * nx;
    a.N = N; a.batch = batch; a.bpc = (uint32_t)p.bpc; a.max_iter = cfg.max_iter; a.two_buffers = (p.nbuf == 2); a.use_tma = 0;
    if (p.fused_kernel) {
        // bulk copies (TMA) need 16-byte aligned sources (strides are multiples of 16 for even nx)
        // (odd nx: a block-row is not a multiple of 16 bytes, so a CTA's share of the matrix is not bulk-copyable)
        const bool aligned = ((uintptr_t)S % 16 == 0) && (!P || (uintptr_t)P % 16 == 0) && ((a.mat_stride * sizeof(T)) % 16 == 0) &&
                             ((3ull * nx * nx * sizeof(T)) % 16 == 0);
        a.use_tma = (aligned && p.pmode != P_SMEM && !(cfg.flags & GBDPCG_FLAG_NO_TMA)) ? 1u : 0u;
        if (p.pmode == P_TMEM && !a.use_tma) {
            if (cfg.flags & GBDPCG_FLAG_TMEM) return fail(GBDPCG_ERR_UNSUPPORTED, "GBDPCG_FLAG_TMEM needs 16-byte aligned S and Pinv");
            // chosen by the planner: fall back to the register-row plan, which can stage with ordinary loads
            gbdpcg_config c2 = cfg;
            c2.flags |= GBDPCG_FLAG_NO_TMEM;
            Plan p2;
            if ((rc = make_plan<T>(nx, N, P != nullptr, c2, batch, &p2))) return rc;
            return launch<T>(p2, S, P, gamma, lambda, nx, N, batch, c2, d_iters, d_exits, stream);
        }
    }
    a.tol = (T)cfg.exit_tol;

    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(n_groups * (uint32_t)p.cluster);
    lc.blockDim = dim3(p.threads);
    lc.dynamicSmemBytes = p.smem;
    lc.stream = stream;
    cudaLaunchAttribute at[1];
    if (p.cluster > 1) {
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = p.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        lc.attrs = at; lc.numAttrs = 1;
    }
    if (p.generic) {
        uint32_t nxv = nx;
        void *args[] = {&a, &nxv};
        CK(cudaLaunchKernelExC(&lc, p.fn, args));
    } else {
        void *args[] = {&a};
        CK(cudaLaunchKernelExC(&lc, p.fn, args));
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

static int check_args(const void *S, const void *gamma, const void *lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg)
{
    if (!S || !gamma || !lambda) return fail(GBDPCG_ERR_INVALID, "S, gamma and lambda must not be NULL");
    if (!cfg) return fail(GBDPCG_ERR_INVALID, "config must not be NULL");
    if (nx == 0) return fail(GBDPCG_ERR_INVALID, "state size must be > 0");
    if (N < 2) return fail(GBDPCG_ERR_INVALID, "knot points must be >= 2 (the reference reads block 1 unconditionally, utils.cuh:21-25)");
    if (cfg->cluster && cfg->cluster != 1 && cfg->cluster != 2 && cfg->cluster != 4 && cfg->cluster != 8 && cfg->cluster != 16)
        return fail(GBDPCG_ERR_INVALID, "config.cluster must be 0,1,2,4,8 or 16");
    if ((uint64_t)nx * N > (1u << 24)) return fail(GBDPCG_ERR_UNSUPPORTED, "system too large");
    return GBDPCG_OK;
}

}  // namespace gbdpcg

// ------------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------------
struct gbdpcg_workspace {
    int device = -1;
    std::mutex mu;
    cudaStream_t streams[2] = {nullptr, nullptr};
    void *dbuf[8] = {};    // S, Pinv, gamma, lambda, then CSR staging: row_ptr, col_ind, val, status
    size_t dcap[8] = {};
    uint32_t *d_iters = nullptr;
    uint8_t *d_exits = nullptr;
    uint32_t *p_iters = nullptr;   // pinned host mirrors of the per-problem flags: a D2H copy into the caller's (pageable)
    uint8_t *p_exits = nullptr;    // arrays would block the host inside the chunk loop and serialise copies and solves
    size_t cap_batch = 0;
    // mapped pinned scalars for the blocking single-solve calls: the kernel writes them in place
    uint32_t *h_iters = nullptr;
    uint8_t *h_exit = nullptr;
    uint32_t *hd_iters = nullptr;
    uint8_t *hd_exit = nullptr;
};

namespace gbdpcg {

static int ws_init(gbdpcg_workspace *w)
{
    CK(cudaGetDevice(&w->device));
    for (auto &s : w->streams) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    void *h = nullptr;
    CK(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
    w->h_iters = (uint32_t *)h;
    w->h_exit = (uint8_t *)h + 32;
    void *d = nullptr;
    CK(cudaHostGetDevicePointer(&d, h, 0));
    w->hd_iters = (uint32_t *)d;
    w->hd_exit = (uint8_t *)d + 32;
    return GBDPCG_OK;
}

static int ws_reserve(gbdpcg_workspace *w, int which, size_t bytes)
{
    if (w->dcap[which] >= bytes) return GBDPCG_OK;
    if (w->dbuf[which]) { CK(cudaFree(w->dbuf[which])); w->dbuf[which] = nullptr; w->dcap[which] = 0; }
    size_t cap = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&w->dbuf[which], cap);
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return fail(GBDPCG_ERR_NOMEM, "cudaMalloc of %zu bytes failed", cap); }
    if (e != cudaSuccess) return fail_cuda(e, "cudaMalloc", __LINE__);
    w->dcap[which] = cap;
    return GBDPCG_OK;
}

static int ws_reserve_flags(gbdpcg_workspace *w, size_t batch)
{
    if (w->cap_batch >= batch) return GBDPCG_OK;
    if (w->d_iters) { CK(cudaFree(w->d_iters)); w->d_iters = nullptr; }
    if (w->d_exits) { CK(cudaFree(w->d_exits)); w->d_exits = nullptr; }
    if (w->p_iters) { CK(cudaFreeHost(w->p_iters)); w->p_iters = nullptr; }
    if (w->p_exits) { CK(cudaFreeHost(w->p_exits)); w->p_exits = nullptr; }
    w->cap_batch = 0;
    size_t cap = batch + batch / 4 + 64;
    CK(cudaMalloc(&w->d_iters, cap * sizeof(uint32_t)));
    CK(cudaMalloc(&w->d_exits, cap));
    CK(cudaHostAlloc((void **)&w->p_iters, cap * sizeof(uint32_t), cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&w->p_exits, cap, cudaHostAllocDefault));
    w->cap_batch = cap;
    return GBDPCG_OK;
}

static void ws_release(gbdpcg_workspace *w)
{
    for (auto &b : w->dbuf) if (b) cudaFree(b);
    if (w->d_iters) cudaFree(w->d_iters);
    if (w->d_exits) cudaFree(w->d_exits);
    if (w->p_iters) cudaFreeHost(w->p_iters);
    if (w->p_exits) cudaFreeHost(w->p_exits);
    if (w->h_iters) cudaFreeHost(w->h_iters);
    for (auto &s : w->streams) if (s) cudaStreamDestroy(s);
}

// process-wide default workspaces, one per device
static std::mutex g_ws_mu;
static std::vector<gbdpcg_workspace *> g_default_ws;

static int get_ws(gbdpcg_workspace *user, gbdpcg_workspace **out)
{
    if (user) { *out = user; return GBDPCG_OK; }
    int dev;
    CK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_ws_mu);
    for (auto *w : g_default_ws)
        if (w->device == dev) { *out = w; return GBDPCG_OK; }
    auto *w = new gbdpcg_workspace();
    int rc = ws_init(w);
    if (rc) { ws_release(w); delete w; return rc; }
    g_default_ws.push_back(w);
    *out = w;
    return GBDPCG_OK;
}

// Blocking host entry points: a failure after the first enqueue must not return while copies into the caller's arrays or kernels on
// the workspace buffers are still in flight.  Synchronises the stream on scope exit unless disarmed after the regular synchronize.
struct StreamDrain {
    cudaStream_t st;
    bool armed = true;
    ~StreamDrain() { if (armed) { cudaStreamSynchronize(st); cudaGetLastError(); } }
};

// ------------------------------------------------------------------------------------------------
// typed implementations
// ------------------------------------------------------------------------------------------------
static inline bool use_pinv(const void *Pinv, const gbdpcg_config *cfg) { return Pinv != nullptr && cfg->empty_pinv == 0; }

// GBDPCG_FLAG_BUILD_*: the host entry points form Phi^-1 from S on the device instead of taking it from the caller
static inline int build_kind(const gbdpcg_config *cfg)
{
    if (cfg->flags & GBDPCG_FLAG_BUILD_STAIR) return GBDPCG_PINV_STAIR;
    if (cfg->flags & GBDPCG_FLAG_BUILD_JACOBI) return GBDPCG_PINV_JACOBI;
    return -1;
}

template <typename T>
static int build_pinv_dev(const T *S, T *Pinv, uint32_t nx, uint32_t N, uint32_t batch, int kind, void *stream)
{
    if (!S || !Pinv) return fail(GBDPCG_ERR_INVALID, "S and Pinv must not be NULL");
    if (nx == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0 and N >= 2");
    if (kind != GBDPCG_PINV_IDENTITY && kind != GBDPCG_PINV_JACOBI && kind != GBDPCG_PINV_STAIR)
        return fail(GBDPCG_ERR_INVALID, "preconditioner kind must be GBDPCG_PINV_IDENTITY, _JACOBI or _STAIR");
    if ((uint64_t)nx * N > (1u << 24)) return fail(GBDPCG_ERR_UNSUPPORTED, "system too large");
    if (batch == 0) return GBDPCG_OK;
    const int rc = launch_pinv_build<T>(S, Pinv, nx, N, batch, kind, (cudaStream_t)stream, nullptr, false);
    if (rc == -3) return fail(GBDPCG_ERR_UNSUPPORTED, "state size %u is too large for the preconditioner builder", nx);
    if (rc != 0) return fail_cuda((cudaError_t)rc, "launch_pinv_build", __LINE__);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

template <typename T>
static int build_pinv_host(gbdpcg_workspace *user_ws, const T *hS, T *hP, uint32_t nx, uint32_t N, uint32_t batch, int kind);

// H2D of `n` systems of S.  GBDPCG_FLAG_SYMMETRIC_S: only the diagonal and right blocks cross PCIe (one strided copy, 2/3 of the
// bytes); the left blocks are rebuilt on the device as transposes of the right blocks of the block-row above (sym_fill.cuh).
template <typename T>
static int upload_S(T *dS, const T *hS, uint32_t nx, uint32_t N, size_t n, const gbdpcg_config *cfg, cudaStream_t st)
{
    const size_t bb = (size_t)nx * nx;
    if (!(cfg->flags & GBDPCG_FLAG_SYMMETRIC_S)) {
        CK(cudaMemcpyAsync(dS, hS, sizeof(T) * 3 * bb * N * n, cudaMemcpyHostToDevice, st));
        return GBDPCG_OK;
    }
    CK(cudaMemcpy2DAsync(dS + bb, sizeof(T) * 3 * bb, hS + bb, sizeof(T) * 3 * bb, sizeof(T) * 2 * bb, (size_t)N * n, cudaMemcpyHostToDevice, st));
    const unsigned long long items = (unsigned long long)N * n;
    const unsigned long long cap = (unsigned long long)device_sms() * 8;
    sym_fill_left_kernel<T><<<(unsigned)(items < cap ? items : cap), 256, 0, st>>>(dS, nx, N, (uint32_t)n);
    CK(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

template <typename T>
static int solve_batched_dev(const T *S, const T *Pinv, const T *gamma, T *lambda, uint32_t nx, uint32_t N, uint32_t batch,
                             const gbdpcg_config *cfg, uint32_t *d_iters, uint8_t *d_exits, void *stream)
{
    int rc = check_args(S, gamma, lambda, nx, N, cfg);
    if (rc) return rc;
    const bool has_p = use_pinv(Pinv, cfg);
    Plan p;
    rc = make_plan<T>(nx, N, has_p, *cfg, batch, &p);
    if (rc) return rc;
    return launch<T>(p, S, has_p ? Pinv : nullptr, gamma, lambda, nx, N, batch, *cfg, d_iters, d_exits, (cudaStream_t)stream);
}

template <typename T>
static int solve_dev_sync(gbdpcg_workspace *user_ws, const T *S, const T *Pinv, const T *gamma, T *lambda, uint32_t nx, uint32_t N,
                          const gbdpcg_config *cfg, uint32_t *iters, int32_t *max_iter_exit, void *stream)
{
    gbdpcg_workspace *w;
    int rc = get_ws(user_ws, &w);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(w->mu);
    rc = solve_batched_dev<T>(S, Pinv, gamma, lambda, nx, N, 1, cfg, w->hd_iters, w->hd_exit, stream);
    if (rc) return rc;
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    if (iters) *iters = *(volatile uint32_t *)w->h_iters;
    if (max_iter_exit) *max_iter_exit = *(volatile uint8_t *)w->h_exit;
    return GBDPCG_OK;
}

template <typename T>
static int solve_host(gbdpcg_workspace *user_ws, const T *hS, const T *hP, const T *hg, T *hl, uint32_t nx, uint32_t N,
                      const gbdpcg_config *cfg, uint32_t *iters, int32_t *max_iter_exit)
{
    int rc = check_args(hS, hg, hl, nx, N, cfg);
    if (rc) return rc;
    gbdpcg_workspace *w;
    rc = get_ws(user_ws, &w);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(w->mu);
    const int bk = build_kind(cfg);
    const bool has_p = bk >= 0 || use_pinv(hP, cfg);
    const size_t mb = sizeof(T) * 3 * (size_t)N * nx * nx, vb = sizeof(T) * (size_t)N * nx;
    if ((rc = ws_reserve(w, 0, mb))) return rc;
    if (has_p && (rc = ws_reserve(w, 1, mb))) return rc;
    if ((rc = ws_reserve(w, 2, vb))) return rc;
    if ((rc = ws_reserve(w, 3, vb))) return rc;
    cudaStream_t st = w->streams[0];
    StreamDrain drain{st};
    if ((rc = upload_S<T>((T *)w->dbuf[0], hS, nx, N, 1, cfg, st))) return rc;
    if (bk >= 0) { if ((rc = build_pinv_dev<T>((const T *)w->dbuf[0], (T *)w->dbuf[1], nx, N, 1, bk, st))) return rc; }
    else if (has_p) CK(cudaMemcpyAsync(w->dbuf[1], hP, mb, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(w->dbuf[2], hg, vb, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(w->dbuf[3], hl, vb, cudaMemcpyHostToDevice, st));
    gbdpcg_config c1 = *cfg;
    if (bk >= 0) c1.empty_pinv = 0;
    rc = solve_batched_dev<T>((const T *)w->dbuf[0], has_p ? (const T *)w->dbuf[1] : nullptr, (const T *)w->dbuf[2], (T *)w->dbuf[3],
                              nx, N, 1, &c1, w->hd_iters, w->hd_exit, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(hl, w->dbuf[3], vb, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    drain.armed = false;
    if (iters) *iters = *(volatile uint32_t *)w->h_iters;
    if (max_iter_exit) *max_iter_exit = *(volatile uint8_t *)w->h_exit;
    return GBDPCG_OK;
}

// Host-resident batch: the batch is cut into chunks that alternate between two streams, so the H2D copy
// of chunk c+1 and the D2H copy of chunk c-1 overlap the solve of chunk c (with pinned host memory;
// pageable memory still works, the copies then serialise inside the driver).
template <typename T>
static int solve_batched_host(gbdpcg_workspace *user_ws, const T *hS, const T *hP, const T *hg, T *hl, uint32_t nx, uint32_t N,
                              uint32_t batch, const gbdpcg_config *cfg, uint32_t *h_iters, uint8_t *h_exits)
{
    int rc = check_args(hS, hg, hl, nx, N, cfg);
    if (rc) return rc;
    if (batch == 0) return GBDPCG_OK;
    gbdpcg_workspace *w;
    rc = get_ws(user_ws, &w);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(w->mu);
    const int bk = build_kind(cfg);
    const bool has_p = bk >= 0 || use_pinv(hP, cfg);
    const size_t me = 3 * (size_t)N * nx * nx, ve = (size_t)N * nx;
    if ((rc = ws_reserve(w, 0, sizeof(T) * me * batch))) return rc;
    if (has_p && (rc = ws_reserve(w, 1, sizeof(T) * me * batch))) return rc;
    if ((rc = ws_reserve(w, 2, sizeof(T) * ve * batch))) return rc;
    if ((rc = ws_reserve(w, 3, sizeof(T) * ve * batch))) return rc;
    if ((rc = ws_reserve_flags(w, batch))) return rc;

    Plan p;
    rc = make_plan<T>(nx, N, has_p, *cfg, batch, &p);
    if (rc) return rc;
    if ((rc = prepare_fn(p.fn, p.cluster > 1 && !p.grid_kernel))) return rc;
    int resident = 1;
    if ((rc = max_resident_problems(p, &resident))) return rc;
    if (resident < 1) resident = 1;
    // chunk = a whole number of resident waves, about 1/8 of the batch, at least 2 waves
    size_t parts = 8;
    if (const char *e = getenv("GBDPCG_HOST_CHUNKS")) { const long v = atol(e); if (v >= 1 && v <= 1024) parts = (size_t)v; }   // tuning knob
    size_t chunk = ((size_t)batch / parts + resident - 1) / resident * resident;
    if (chunk < 2 * (size_t)resident) chunk = 2 * (size_t)resident;
    if (chunk > batch) chunk = batch;

    T *dS = (T *)w->dbuf[0], *dP = (T *)w->dbuf[1], *dg = (T *)w->dbuf[2], *dl = (T *)w->dbuf[3];
    // A failure after the first enqueue must not return while earlier chunks are still in flight (D2H copies into the caller's
    // lambda, kernels on the workspace buffers): drain both streams first, ignoring what the drain itself reports.
    struct Drain {
        gbdpcg_workspace *w; bool armed = true;
        ~Drain() { if (armed) { cudaStreamSynchronize(w->streams[0]); cudaStreamSynchronize(w->streams[1]); cudaGetLastError(); } }
    } drain{w};
    // The run is H2D-bound: what the caller waits for after the last byte has crossed PCIe is the LAST piece's builder + solve + D2H, so the
    // last chunk is cut in two (down to one resident wave) -- its first half computes while the second half is still arriving.
    std::vector<std::pair<size_t, size_t>> pieces;
    for (size_t o = 0; o < batch; o += chunk) pieces.push_back({o, (batch - o < chunk) ? batch - o : chunk});
    if (pieces.back().second >= 2 * (size_t)resident) {
        const std::pair<size_t, size_t> last = pieces.back();
        const size_t tail = (last.second / 2) / (size_t)resident * (size_t)resident;      // whole waves, at least one (last.second >= 2 waves)
        pieces.back() = {last.first, last.second - tail};
        pieces.push_back({last.first + last.second - tail, tail});
    }
    int ci = 0;
    for (const auto &pc : pieces) {
        const size_t o = pc.first, n = pc.second;
        cudaStream_t st = w->streams[ci++ & 1];
        // every H2D copy of the chunk before its first kernel: a copy queued behind the chunk's own sym_fill / builder kernels would sit at
        // the head of the copy engine's queue until they finish and hold back the next chunk's S
        CK(cudaMemcpyAsync(dg + o * ve, hg + o * ve, sizeof(T) * ve * n, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(dl + o * ve, hl + o * ve, sizeof(T) * ve * n, cudaMemcpyHostToDevice, st));
        if (bk < 0 && has_p) CK(cudaMemcpyAsync(dP + o * me, hP + o * me, sizeof(T) * me * n, cudaMemcpyHostToDevice, st));
        if ((rc = upload_S<T>(dS + o * me, hS + o * me, nx, N, n, cfg, st))) return rc;
        if (bk >= 0) { if ((rc = build_pinv_dev<T>(dS + o * me, dP + o * me, nx, N, (uint32_t)n, bk, st))) return rc; }
        rc = launch<T>(p, dS + o * me, has_p ? dP + o * me : nullptr, dg + o * ve, dl + o * ve, nx, N, (uint32_t)n, *cfg,
                       w->d_iters + o, w->d_exits + o, st);
        if (rc) return rc;
        CK(cudaMemcpyAsync(hl + o * ve, dl + o * ve, sizeof(T) * ve * n, cudaMemcpyDeviceToHost, st));
        if (h_iters) CK(cudaMemcpyAsync(w->p_iters + o, w->d_iters + o, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
        if (h_exits) CK(cudaMemcpyAsync(w->p_exits + o, w->d_exits + o, n, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(w->streams[0]));
    CK(cudaStreamSynchronize(w->streams[1]));
    drain.armed = false;
    if (h_iters) memcpy(h_iters, w->p_iters, sizeof(uint32_t) * batch);
    if (h_exits) memcpy(h_exits, w->p_exits, batch);
    return GBDPCG_OK;
}

template <typename T>
static int build_pinv_host(gbdpcg_workspace *user_ws, const T *hS, T *hP, uint32_t nx, uint32_t N, uint32_t batch, int kind)
{
    if (!hS || !hP) return fail(GBDPCG_ERR_INVALID, "S and Pinv must not be NULL");
    if (nx == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0 and N >= 2");
    if (batch == 0) return GBDPCG_OK;
    gbdpcg_workspace *w;
    int rc = get_ws(user_ws, &w);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(w->mu);
    const size_t mb = sizeof(T) * 3 * (size_t)N * nx * nx * batch;
    if ((rc = ws_reserve(w, 0, mb))) return rc;
    if ((rc = ws_reserve(w, 1, mb))) return rc;
    cudaStream_t st = w->streams[0];
    StreamDrain drain{st};
    CK(cudaMemcpyAsync(w->dbuf[0], hS, mb, cudaMemcpyHostToDevice, st));
    if ((rc = build_pinv_dev<T>((const T *)w->dbuf[0], (T *)w->dbuf[1], nx, N, batch, kind, st))) return rc;
    CK(cudaMemcpyAsync(hP, w->dbuf[1], mb, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    drain.armed = false;
    return GBDPCG_OK;
}

// ---- Schur-complement formation / primal recovery (the steps either side of the solve; schur_form.cuh) ----
template <typename T>
static int schur_form_dev(const T *Q, const T *R, const T *A, const T *B, const T *q, const T *r, const T *c, uint32_t nx, uint32_t nu,
                          uint32_t N, uint32_t batch, T *S, T *gamma, T *Qinv, T *Rinv, cudaStream_t st)
{
    if (!Q || !R || !A || !B || !q || !r || !c || !S || !gamma) return fail(GBDPCG_ERR_INVALID, "stage data, S and gamma must not be NULL");
    if (nx == 0 || nu == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0, nu > 0 and N >= 2");
    if ((uint64_t)nx * N > (1u << 24)) return fail(GBDPCG_ERR_UNSUPPORTED, "system too large");
    if (batch == 0) return GBDPCG_OK;
    SchurArgs a;
    a.Q = Q; a.R = R; a.A = A; a.B = B; a.q = q; a.r = r; a.c = c; a.S = S; a.gamma = gamma; a.Qinv = Qinv; a.Rinv = Rinv;
    a.nx = nx; a.nu = nu; a.N = N; a.batch = batch;
    const int sms = device_sms();
    // compiled state sizes with nu <= nx: chunked, lane-per-columns kernel
    const void *fn_nx = nullptr;
    if (nu <= nx) {
        switch (nx) {
#define CASE(NX) case NX: fn_nx = (const void *)schur_form_nx_kernel<T, NX>; break;
            GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
            default: break;
        }
    }
    if (fn_nx) {
        const int threads = 256;
        int chunk = 0;
        for (size_t budget : {(size_t)112 * 1024, kMaxDynSmem}) {
            int c = (int)N;
            while (c > 1 && schur_nx_smem_bytes<T>((int)nx, c, threads) > budget) c = (c + 1) / 2;
            if (schur_nx_smem_bytes<T>((int)nx, c, threads) <= budget) { chunk = c; break; }
        }
        if (chunk > 0) {
            if (chunk > 4 && (unsigned long long)batch * ((N + chunk - 1) / chunk) < (unsigned long long)sms) {
                unsigned long long want = ((unsigned long long)batch * N + 2ull * sms - 1) / (2ull * sms);
                int c = (int)(want < 4 ? 4 : want);
                if (c < chunk) chunk = c;
            }
            const uint32_t nchunks = (N + chunk - 1) / chunk;
            const size_t smem = schur_nx_smem_bytes<T>((int)nx, chunk, threads);
            int rc = prepare_fn(fn_nx, false);
            if (rc) return rc;
            const unsigned long long items = (unsigned long long)batch * nchunks;
            const int per_sm = smem <= (size_t)112 * 1024 ? 2 : 1;
            const uint32_t grid = (uint32_t)(items < (unsigned long long)(sms * per_sm) ? items : (unsigned long long)(sms * per_sm));
            uint32_t chunk_u = (uint32_t)chunk, nchunks_u = nchunks;
            void *args[] = {&a, &chunk_u, &nchunks_u};
            CK(cudaLaunchKernel(fn_nx, dim3(grid), dim3(threads), args, smem, st));
            g_launches.fetch_add(1, std::memory_order_relaxed);
            return GBDPCG_OK;
        }
    }
    const size_t smem = schur_form_smem_elems(nx, nu) * sizeof(T) + 16;
    if (smem > kMaxDynSmem) return fail(GBDPCG_ERR_UNSUPPORTED, "nx=%u nu=%u needs %zu B of shared memory (limit %zu)", nx, nu, smem, kMaxDynSmem);
    const void *fn = (const void *)schur_form_kernel<T>;
    int rc = prepare_fn(fn, false);
    if (rc) return rc;
    const unsigned long long items = (unsigned long long)batch * N;
    const unsigned long long cap16 = (unsigned long long)sms * 16;
    const uint32_t grid = (uint32_t)(items < cap16 ? items : cap16);
    schur_form_kernel<T><<<grid, 128, smem, st>>>(a);
    CK(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

template <typename T>
static int schur_recover_dev(const T *Qinv, const T *Rinv, const T *A, const T *B, const T *q, const T *r, const T *lambda, uint32_t nx,
                             uint32_t nu, uint32_t N, uint32_t batch, T *x, T *u, cudaStream_t st)
{
    if (!Qinv || !Rinv || !A || !B || !q || !r || !lambda || !x || !u) return fail(GBDPCG_ERR_INVALID, "arguments must not be NULL");
    if (nx == 0 || nu == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0, nu > 0 and N >= 2");
    if (batch == 0) return GBDPCG_OK;
    const unsigned long long items = (unsigned long long)batch * N;
    const unsigned long long sms = (unsigned long long)device_sms();
    const size_t staged_bytes = (size_t)8 * 3 * schur_recover_slice_elems<T>(nx, nu) * sizeof(T);
    const char *env_rec = getenv("GBDPCG_RECOVER_GLOBAL");                     // A/B knob: the global-memory form of the warp kernel
    const bool staged = !(env_rec && atoi(env_rec));
    if (nx <= 32 && nu <= 32 && staged_bytes <= 96 * 1024 && staged) {
        // one warp per (system, knot), the knot's matrices staged through the warp's slice of shared memory (two slices: next item in flight)
        void (*fn)(const T *, const T *, const T *, const T *, const T *, const T *, const T *, T *, T *, uint32_t, uint32_t, uint32_t, uint32_t) =
            schur_recover_staged_kernel<T, 0, 0>;
        if (2 * nu == nx) {                                                    // (the usual shape: half as many inputs as states)
            switch (nx) {
#define CASE(NX) case NX: fn = schur_recover_staged_kernel<T, NX, NX / 2>; break;
                GBDPCG_FOR_EACH_NX(CASE)
#undef CASE
                default: break;
            }
        }
        int rc;
        if ((rc = prepare_fn(reinterpret_cast<const void *>(fn), false))) return rc;
        const unsigned long long per_sm = std::min<unsigned long long>(sizeof(T) == 4 ? 4 : 2, std::max<unsigned long long>(1, (227 * 1024) / (staged_bytes + 1024)));
        const unsigned long long ctas = (items + 7) / 8;
        const uint32_t grid = (uint32_t)std::min(ctas, sms * per_sm);
        fn<<<grid, 256, staged_bytes, st>>>(Qinv, Rinv, A, B, q, r, lambda, x, u, nx, nu, N, batch);
    } else if (nx <= 32 && nu <= 32) {
        const unsigned long long ctas = (items + 7) / 8;                       // one warp per (system, knot)
        const uint32_t grid = (uint32_t)(ctas < sms * 8 ? ctas : sms * 8);
        schur_recover_warp_kernel<T><<<grid, 256, 0, st>>>(Qinv, Rinv, A, B, q, r, lambda, x, u, nx, nu, N, batch);
    } else {
        const uint32_t grid = (uint32_t)(items < sms * 32 ? items : sms * 32);
        schur_recover_kernel<T><<<grid, 64, (nx + nu) * sizeof(T) + 16, st>>>(Qinv, Rinv, A, B, q, r, lambda, x, u, nx, nu, N, batch);
    }
    CK(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

// ---- CSR front end (reference include/interface.cuh:8-20 is a stub) ----
template <typename T>
static int pack_csr_dev(const uint32_t *row_ptr, const uint32_t *col_ind, const T *val, uint32_t nnz, uint32_t rows, uint32_t nx, uint32_t N,
                        T *out, int32_t *d_status, cudaStream_t st)
{
    if (!row_ptr || !out || (nnz && (!col_ind || !val))) return fail(GBDPCG_ERR_INVALID, "CSR arrays and the output must not be NULL");
    if (nx == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0 and N >= 2");
    if ((uint64_t)rows != (uint64_t)nx * N) return fail(GBDPCG_ERR_INVALID, "CSR matrix has %u rows, expected stateSize*knotPoints = %llu", rows,
                                                        (unsigned long long)nx * N);
    CK(cudaMemsetAsync(out, 0, sizeof(T) * 3 * (size_t)N * nx * nx, st));
    if (d_status) CK(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    const uint32_t cap8 = (uint32_t)device_sms() * 8;
    const uint32_t blocks = (rows + 7) / 8 < cap8 ? (rows + 7) / 8 : cap8;       // one warp per row, at most 8 CTAs per SM
    csr_pack_kernel<T><<<blocks, 256, 0, st>>>(row_ptr, col_ind, val, nnz, rows, nx, N, out, d_status);
    CK(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

struct HostCsr { const uint32_t *row_ptr, *col_ind; const void *val; uint32_t nnz; };

template <typename T>
static int solve_csr_host(gbdpcg_workspace *user_ws, HostCsr S, HostCsr P, const T *hg, T *hl, uint32_t nx, uint32_t N,
                          const gbdpcg_config *cfg, uint32_t *iters, int32_t *max_iter_exit)
{
    int rc = check_args(S.val, hg, hl, nx, N, cfg);
    if (rc) return rc;
    if (!S.row_ptr || !S.col_ind) return fail(GBDPCG_ERR_INVALID, "S.row_ptr and S.col_ind must not be NULL");
    // The CSR arrays are host pointers: hold row_ptr to the CSR contract BEFORE anything is uploaded, so a malformed csr_t can
    // neither over-read the device staging buffers (sized from nnz) nor fault the context.
    {
        const uint32_t nrows = nx * N;
        auto check_csr = [&](const HostCsr &M, const char *name) -> int {
            if (M.row_ptr[0] != 0) return fail(GBDPCG_ERR_INVALID, "%s.row_ptr[0] must be 0 (is %u)", name, M.row_ptr[0]);
            for (uint32_t i = 0; i < nrows; i++)
                if (M.row_ptr[i + 1] < M.row_ptr[i]) return fail(GBDPCG_ERR_INVALID, "%s.row_ptr decreases at row %u", name, i);
            if (M.row_ptr[nrows] > M.nnz)
                return fail(GBDPCG_ERR_INVALID, "%s.row_ptr[rows] = %u exceeds nnz = %u", name, M.row_ptr[nrows], M.nnz);
            return GBDPCG_OK;
        };
        if ((rc = check_csr(S, "S"))) return rc;
        if (P.val && P.row_ptr && P.col_ind && cfg->empty_pinv == 0 && build_kind(cfg) < 0 && (rc = check_csr(P, "Pinv"))) return rc;
    }
    gbdpcg_workspace *w;
    if ((rc = get_ws(user_ws, &w))) return rc;
    std::lock_guard<std::mutex> lk(w->mu);
    const int bk = build_kind(cfg);
    const bool csr_p = P.val && P.row_ptr && P.col_ind && cfg->empty_pinv == 0 && bk < 0;
    const bool has_p = bk >= 0 || csr_p;
    const uint32_t rows = nx * N;
    const size_t mb = sizeof(T) * 3 * (size_t)N * nx * nx, vb = sizeof(T) * (size_t)N * nx;
    const uint32_t max_nnz = (csr_p && P.nnz > S.nnz) ? P.nnz : S.nnz;
    if ((rc = ws_reserve(w, 0, mb))) return rc;
    if (has_p && (rc = ws_reserve(w, 1, mb))) return rc;
    if ((rc = ws_reserve(w, 2, vb))) return rc;
    if ((rc = ws_reserve(w, 3, vb))) return rc;
    if ((rc = ws_reserve(w, 4, sizeof(uint32_t) * ((size_t)rows + 1)))) return rc;
    if ((rc = ws_reserve(w, 5, sizeof(uint32_t) * (size_t)(max_nnz ? max_nnz : 1)))) return rc;
    if ((rc = ws_reserve(w, 6, sizeof(T) * (size_t)(max_nnz ? max_nnz : 1)))) return rc;
    if ((rc = ws_reserve(w, 7, 2 * sizeof(int32_t)))) return rc;
    cudaStream_t st = w->streams[0];
    StreamDrain drain{st};
    int32_t *d_status = (int32_t *)w->dbuf[7];
    auto upload_and_pack = [&](const HostCsr &M, T *dst, int32_t *stat) -> int {
        CK(cudaMemcpyAsync(w->dbuf[4], M.row_ptr, sizeof(uint32_t) * ((size_t)rows + 1), cudaMemcpyHostToDevice, st));
        if (M.nnz) {
            CK(cudaMemcpyAsync(w->dbuf[5], M.col_ind, sizeof(uint32_t) * (size_t)M.nnz, cudaMemcpyHostToDevice, st));
            CK(cudaMemcpyAsync(w->dbuf[6], M.val, sizeof(T) * (size_t)M.nnz, cudaMemcpyHostToDevice, st));
        }
        return pack_csr_dev<T>((const uint32_t *)w->dbuf[4], (const uint32_t *)w->dbuf[5], (const T *)w->dbuf[6], M.nnz, rows, nx, N, dst, stat, st);
    };
    if ((rc = upload_and_pack(S, (T *)w->dbuf[0], d_status))) return rc;
    if (csr_p) { if ((rc = upload_and_pack(P, (T *)w->dbuf[1], d_status + 1))) return rc; }
    else if (bk >= 0) { if ((rc = build_pinv_dev<T>((const T *)w->dbuf[0], (T *)w->dbuf[1], nx, N, 1, bk, st))) return rc; }
    // the band check must pass before the solve is worth running
    int32_t h_status[2] = {0, 0};
    CK(cudaMemcpyAsync(h_status, d_status, (csr_p ? 2 : 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(w->dbuf[2], hg, vb, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(w->dbuf[3], hl, vb, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    if (h_status[0] || h_status[1])
        return fail(GBDPCG_ERR_INVALID, "%s has entries outside the block-tridiagonal band of block size %u (or a malformed row_ptr)",
                    h_status[0] ? "S" : "Pinv", nx);
    gbdpcg_config c1 = *cfg;
    c1.empty_pinv = has_p ? 0 : 1;
    rc = solve_batched_dev<T>((const T *)w->dbuf[0], has_p ? (const T *)w->dbuf[1] : nullptr, (const T *)w->dbuf[2], (T *)w->dbuf[3],
                              nx, N, 1, &c1, w->hd_iters, w->hd_exit, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(hl, w->dbuf[3], vb, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    drain.armed = false;
    if (iters) *iters = *(volatile uint32_t *)w->h_iters;
    if (max_iter_exit) *max_iter_exit = *(volatile uint8_t *)w->h_exit;
    return GBDPCG_OK;
}

template <typename T>
static int plan_query(uint32_t nx, uint32_t N, int has_pinv, uint32_t batch, const gbdpcg_config *cfg, gbdpcg_plan_info *out)
{
    Plan p;
    int rc = make_plan<T>(nx, N, has_pinv != 0, *cfg, batch ? batch : 1u, &p);
    if (rc) return rc;
    memset(out, 0, sizeof *out);
    out->specialised = p.generic ? 0 : 1;
    out->cluster = p.cluster; out->threads = p.threads; out->block_rows = p.bpc; out->pinv_mode = p.pmode;
    out->smem_bytes = p.smem;
    snprintf(out->kernel, sizeof out->kernel, "%s", p.name);
    out->regs = -1; out->max_resident = -1;
    // device-dependent facts: best effort (absent on a machine without a GPU)
    cudaFuncAttributes fa;
    if (cudaFuncGetAttributes(&fa, p.fn) == cudaSuccess) out->regs = fa.numRegs; else cudaGetLastError();
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) == cudaSuccess && dev_count > 0) {
        int res = 0;
        if (prepare_fn(p.fn, p.cluster > 1 && !p.grid_kernel, p.ctas_per_sm > 1) == GBDPCG_OK && max_resident_problems(p, &res) == GBDPCG_OK) out->max_resident = res;
    } else {
        cudaGetLastError();
    }
    return GBDPCG_OK;
}

}  // namespace gbdpcg

// ------------------------------------------------------------------------------------------------
// allocation-free solve (gbdpcg_solve_dev_static_scratch): the cooperative-grid kernel with its exchange scratch in a
// file-scope device array instead of the stream-ordered pool allocation of launch() above
// ------------------------------------------------------------------------------------------------
// Every GPU that loads the library holds these GBDPCG_STATIC_SCRATCH_BYTES until exit.  In host code the name below is only a
// host-side stand-in for the symbol: the kernel gets the device address that cudaGetSymbolAddress returns (static_scratch_ptr).
static __device__ __align__(256) unsigned char g_static_scratch[GBDPCG_STATIC_SCRATCH_BYTES];

namespace gbdpcg {

// device address of g_static_scratch on the CURRENT device; looked up once per device, with that device current
static int static_scratch_ptr(void **out)
{
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static void *ptrs[kMaxDev] = {};
    int dev;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) return fail(GBDPCG_ERR_UNSUPPORTED, "device ordinal %d is beyond the %d the static scratch table holds", dev, kMaxDev);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(GBDPCG_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return GBDPCG_OK;
}

// One system on the cooperative-grid kernel (whatever its size: the planner's own path takes this kernel only beyond one cluster).
// No cudaMalloc / cudaFree / pool allocation anywhere below; the kernel overwrites every scratch word before it reads it
// (pcg_grid.cuh: each buffer is written between one pair of grid barriers and read between the next), so nothing has to be cleared.
template <typename T>
static int solve_dev_static_scratch(const T *S, const T *Pinv, const T *gamma, T *lambda, uint32_t nx, uint32_t N, const gbdpcg_config *cfg,
                                    uint32_t *d_iters, uint8_t *d_exits, cudaStream_t stream)
{
    int rc = check_args(S, gamma, lambda, nx, N, cfg);
    if (rc) return rc;
    const bool has_p = use_pinv(Pinv, cfg);
    Plan p;
    // The one-barrier form where its (larger, parity-alternating) exchange buffers fit the static array, else the two-barrier form the
    // array was sized for: the size limit documented in gbdpcg.h -- CTAs * (6 nx + 2) * sizeof(T) -- is unchanged, results are bit-identical.
    bool planned = plan_grid<T>(nx, N, has_p, *cfg, &p, /*allow_one_barrier=*/true);
    if (planned && p.grid_one_barrier && grid1_scratch_elems<T>((int)nx, p.cluster) * sizeof(T) > (size_t)GBDPCG_STATIC_SCRATCH_BYTES)
        planned = plan_grid<T>(nx, N, has_p, *cfg, &p, /*allow_one_barrier=*/false);
    if (!planned)
        return fail(GBDPCG_ERR_UNSUPPORTED, "no cooperative-grid kernel for nx=%u N=%u %s flags=0x%x cluster=%u (even nx <= 16, default recurrence, "
                    "cluster = 0, N within the co-residency limit)", nx, N, sizeof(T) == 8 ? "f64" : "f32", cfg->flags, cfg->cluster);
    const int G = p.cluster;
    const size_t need = (p.grid_one_barrier ? grid1_scratch_elems<T>((int)nx, G) : grid_scratch_elems<T>((int)nx, G)) * sizeof(T);
    if (need > (size_t)GBDPCG_STATIC_SCRATCH_BYTES)
        return fail(GBDPCG_ERR_UNSUPPORTED, "nx=%u N=%u on %d CTAs needs %zu bytes of exchange scratch; the static array holds %d "
                    "(GBDPCG_STATIC_SCRATCH_BYTES) -- use gbdpcg_solve_dev_*", nx, N, G, need, (int)GBDPCG_STATIC_SCRATCH_BYTES);
    void *base = nullptr;
    if ((rc = static_scratch_ptr(&base))) return rc;
    if ((rc = prepare_fn(p.fn, false))) return rc;
    int resident = 0;
    if ((rc = max_resident_problems(p, &resident))) return rc;
    if (resident < 1)
        return fail(GBDPCG_ERR_UNSUPPORTED, "kernel %s cannot become resident (grid=%d threads=%d smem=%zu)", p.name, G, p.threads, p.smem);
    KernelArgs<T> a;
    a.S = S; a.Pinv = has_p ? Pinv : nullptr; a.gamma = gamma; a.lambda = lambda; a.iters = d_iters; a.exits = d_exits;
    a.mat_stride = 3ull * N * nx * nx; a.vec_stride = (unsigned long long)N * nx;
    a.N = N; a.batch = 1; a.bpc = (uint32_t)p.bpc; a.max_iter = cfg->max_iter; a.two_buffers = 0; a.use_tma = 0;
    a.tol = (T)cfg->exit_tol;
    T *scratch = reinterpret_cast<T *>(base);
    GridScratch<T> gs;
    gs.halo_r = scratch; gs.halo_rt = scratch + (size_t)G * 2 * nx; gs.halo_w = scratch + (size_t)G * 4 * nx;
    gs.partials = scratch + (size_t)G * 6 * nx;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)G);
    lc.blockDim = dim3(p.threads);
    lc.dynamicSmemBytes = p.smem;
    lc.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    lc.attrs = at; lc.numAttrs = 1;
    void *args[] = {&a, &gs};
    void *args1[] = {&a, &scratch};
    const cudaError_t e = cudaLaunchKernelExC(&lc, p.fn, p.grid_one_barrier ? args1 : args);
    if (e != cudaSuccess) return fail_cuda(e, "cudaLaunchKernelExC (cooperative grid, static scratch)", __LINE__);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return GBDPCG_OK;
}

}  // namespace gbdpcg

// ------------------------------------------------------------------------------------------------
// extern "C"
// ------------------------------------------------------------------------------------------------
using namespace gbdpcg;

extern "C" {

void gbdpcg_config_default(gbdpcg_config *cfg)
{
    if (!cfg) return;
    cfg->exit_tol = 1e-6;   // reference include/constants.cuh:17
    cfg->max_iter = 25;     // reference include/constants.cuh:15
    cfg->empty_pinv = 1;    // reference include/types.cuh:33
    cfg->flags = 0;
    cfg->cluster = 0;
}

int gbdpcg_workspace_create(gbdpcg_workspace **ws)
{
    if (!ws) return fail(GBDPCG_ERR_INVALID, "ws must not be NULL");
    auto *w = new gbdpcg_workspace();
    int rc = ws_init(w);
    if (rc) { ws_release(w); delete w; *ws = nullptr; return rc; }
    *ws = w;
    return GBDPCG_OK;
}

void gbdpcg_workspace_destroy(gbdpcg_workspace *ws)
{
    if (!ws) return;
    ws_release(ws);
    delete ws;
}

#define DEFINE_TYPED(SUF, T)                                                                                                        \
    int gbdpcg_solve_host_##SUF(gbdpcg_workspace *ws, const T *S, const T *P, const T *g, T *l, uint32_t nx, uint32_t N,            \
                                const gbdpcg_config *cfg, uint32_t *iters, int32_t *mie)                                            \
    { return solve_host<T>(ws, S, P, g, l, nx, N, cfg, iters, mie); }                                                               \
    int gbdpcg_solve_dev_##SUF(gbdpcg_workspace *, const T *S, const T *P, const T *g, T *l, uint32_t nx, uint32_t N,               \
                               const gbdpcg_config *cfg, uint32_t *d_iters, uint8_t *d_mie, void *stream)                           \
    { return solve_batched_dev<T>(S, P, g, l, nx, N, 1, cfg, d_iters, d_mie, stream); }                                             \
    int gbdpcg_solve_dev_sync_##SUF(gbdpcg_workspace *ws, const T *S, const T *P, const T *g, T *l, uint32_t nx, uint32_t N,        \
                                    const gbdpcg_config *cfg, uint32_t *iters, int32_t *mie, void *stream)                          \
    {                                                                                                                               \
        int rc = check_args(S, g, l, nx, N, cfg);                                                                                   \
        return rc ? rc : solve_dev_sync<T>(ws, S, P, g, l, nx, N, cfg, iters, mie, stream);                                         \
    }                                                                                                                               \
    int gbdpcg_solve_batched_dev_##SUF(gbdpcg_workspace *, const T *S, const T *P, const T *g, T *l, uint32_t nx, uint32_t N,       \
                                       uint32_t batch, const gbdpcg_config *cfg, uint32_t *d_iters, uint8_t *d_mie, void *stream)   \
    { return solve_batched_dev<T>(S, P, g, l, nx, N, batch, cfg, d_iters, d_mie, stream); }                                         \
    int gbdpcg_solve_batched_host_##SUF(gbdpcg_workspace *ws, const T *S, const T *P, const T *g, T *l, uint32_t nx, uint32_t N,    \
                                        uint32_t batch, const gbdpcg_config *cfg, uint32_t *h_iters, uint8_t *h_mie)                \
    { return solve_batched_host<T>(ws, S, P, g, l, nx, N, batch, cfg, h_iters, h_mie); }                                            \
    int gbdpcg_build_pinv_dev_##SUF(gbdpcg_workspace *, const T *S, T *P, uint32_t nx, uint32_t N, uint32_t batch, int kind,        \
                                    void *stream)                                                                                   \
    { return build_pinv_dev<T>(S, P, nx, N, batch, kind, stream); }                                                                 \
    int gbdpcg_build_pinv_host_##SUF(gbdpcg_workspace *ws, const T *S, T *P, uint32_t nx, uint32_t N, uint32_t batch, int kind)     \
    { return build_pinv_host<T>(ws, S, P, nx, N, batch, kind); }                                                                   \
    int gbdpcg_schur_form_dev_##SUF(gbdpcg_workspace *, const T *Q, const T *R, const T *A, const T *B, const T *q, const T *r,      \
                                    const T *c, uint32_t nx, uint32_t nu, uint32_t N, uint32_t batch, T *S, T *gamma, T *Qinv,      \
                                    T *Rinv, void *stream)                                                                          \
    { return schur_form_dev<T>(Q, R, A, B, q, r, c, nx, nu, N, batch, S, gamma, Qinv, Rinv, (cudaStream_t)stream); }                \
    int gbdpcg_schur_recover_dev_##SUF(gbdpcg_workspace *, const T *Qinv, const T *Rinv, const T *A, const T *B, const T *q,        \
                                       const T *r, const T *lambda, uint32_t nx, uint32_t nu, uint32_t N, uint32_t batch, T *x,     \
                                       T *u, void *stream)                                                                          \
    { return schur_recover_dev<T>(Qinv, Rinv, A, B, q, r, lambda, nx, nu, N, batch, x, u, (cudaStream_t)stream); }                  \
    int gbdpcg_pack_csr_dev_##SUF(gbdpcg_workspace *, const uint32_t *rp, const uint32_t *ci, const T *v, uint32_t nnz,            \
                                  uint32_t rows, uint32_t nx, uint32_t N, T *out, int32_t *d_status, void *stream)                  \
    { return pack_csr_dev<T>(rp, ci, v, nnz, rows, nx, N, out, d_status, (cudaStream_t)stream); }                                   \
    int gbdpcg_solve_csr_host_##SUF(gbdpcg_workspace *ws, const uint32_t *s_rp, const uint32_t *s_ci, const T *s_v, uint32_t s_nnz, \
                                    const uint32_t *p_rp, const uint32_t *p_ci, const T *p_v, uint32_t p_nnz, const T *g, T *l,     \
                                    uint32_t nx, uint32_t N, const gbdpcg_config *cfg, uint32_t *iters, int32_t *mie)               \
    { return solve_csr_host<T>(ws, HostCsr{s_rp, s_ci, s_v, s_nnz}, HostCsr{p_rp, p_ci, p_v, p_nnz}, g, l, nx, N, cfg, iters, mie); }

DEFINE_TYPED(f32, float)
DEFINE_TYPED(f64, double)
#undef DEFINE_TYPED

int gbdpcg_solve_dev_static_scratch(int is_f64, const void *d_S, const void *d_Pinv, const void *d_gamma, void *d_lambda, uint32_t nx,
                                    uint32_t N, const gbdpcg_config *cfg, uint32_t *d_iters, uint8_t *d_max_iter_exit, void *stream)
{
    if (is_f64)
        return solve_dev_static_scratch<double>((const double *)d_S, (const double *)d_Pinv, (const double *)d_gamma, (double *)d_lambda, nx, N, cfg,
                                                d_iters, d_max_iter_exit, (cudaStream_t)stream);
    return solve_dev_static_scratch<float>((const float *)d_S, (const float *)d_Pinv, (const float *)d_gamma, (float *)d_lambda, nx, N, cfg, d_iters,
                                           d_max_iter_exit, (cudaStream_t)stream);
}

int gbdpcg_plan_query(uint32_t nx, uint32_t N, int is_f64, int has_pinv, const gbdpcg_config *cfg, gbdpcg_plan_info *out)
{
    gbdpcg_config def;
    gbdpcg_config_default(&def);
    if (!cfg) cfg = &def;
    if (!out) return fail(GBDPCG_ERR_INVALID, "out must not be NULL");
    if (nx == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0 and N >= 2");
    return is_f64 ? plan_query<double>(nx, N, has_pinv, 1u, cfg, out) : plan_query<float>(nx, N, has_pinv, 1u, cfg, out);
}

int gbdpcg_plan_query_batched(uint32_t nx, uint32_t N, int is_f64, int has_pinv, uint32_t batch, const gbdpcg_config *cfg,
                              gbdpcg_plan_info *out)
{
    gbdpcg_config def;
    gbdpcg_config_default(&def);
    if (!cfg) cfg = &def;
    if (!out) return fail(GBDPCG_ERR_INVALID, "out must not be NULL");
    if (nx == 0 || N < 2) return fail(GBDPCG_ERR_INVALID, "need nx > 0 and N >= 2");
    return is_f64 ? plan_query<double>(nx, N, has_pinv, batch, cfg, out) : plan_query<float>(nx, N, has_pinv, batch, cfg, out);
}

void *gbdpcg_host_alloc(size_t bytes)
{
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) { fail_cuda(e, "cudaHostAlloc", __LINE__); cudaGetLastError(); return nullptr; }
    return p;
}
void gbdpcg_host_free(void *p) { if (p) cudaFreeHost(p); }

int gbdpcg_version(void) { return GBDPCG_VERSION; }
const char *gbdpcg_build_arch(void) { return "sm_100a"; }
const char *gbdpcg_last_error(void) { return tl_error.c_str(); }
int gbdpcg_last_cuda_error(void) { return tl_cuda_error; }
uint64_t gbdpcg_kernel_launches(void) { return g_launches.load(); }

}  // extern "C"
