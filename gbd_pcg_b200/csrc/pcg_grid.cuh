// pcg_grid.cuh -- systems LARGER than one 16-CTA cluster: a cooperative grid of register-row CTAs owns one system.
//
// The cluster kernel (pcg_fused.cuh) holds a system inside one thread-block cluster: at most 16 CTAs, i.e. N <= 752 knot points for
// nx = 14 in fp32 with register rows, 1024 with Phi^-1 in tensor memory.  The reference has no such limit -- its kernel is one CTA per knot point on a cooperative grid and runs any N
// whose N CTAs are co-resident (reference include/pcg.cuh:23-49, launch at include/interface.cuh:130-132), 4 grid.sync per
// iteration with every vector exchanged through global memory (include/pcg.cuh:154-208).  This kernel covers that range:
//   * cooperative launch of G CTAs (one per SM at most), CTA c owns bpc consecutive block-rows, a thread owns R rows of S and
//     Phi^-1 IN REGISTERS for the whole solve (same ownership and mat-vec as pcg_fused.cuh);
//   * single-reduction (Chronopoulos-Gear) form of the recurrence of include/pcg.cuh:118-212, TWO grid barriers per iteration:
//       r~ = Phi^-1 r            (halo of r comes from GHOST copies kept locally, see below -- no exchange)
//       boundary rows of r~  -> global          -------- grid barrier 1 --------
//       w = S r~ ;  CTA partials of (r.r~, w.r~) and boundary rows of w -> global     -------- grid barrier 2 --------
//       every CTA re-sums the SAME G partials in the SAME order (bit-identical gamma, delta everywhere, hence a uniform exit
//       decision -- the property the reference gets from re-reducing d_eta_new_temp in every block, include/pcg.cuh:189-195);
//       alpha, beta, p, s, lambda, r updates in registers;
//   * ghost rows: a CTA keeps copies of its neighbours' boundary block-rows of r and s; the neighbour's boundary rows of w arrive
//     with the reduction and  s_g = fma(beta, s_g, w_g),  r_g = fma(-alpha, s_g, r_g)  are repeated locally with the same scalars
//     and the same fused multiply-adds as on the owner, so the copies are bit-identical to what an exchange of r would deliver.
// What crosses CTAs goes through a small global scratch (L2) with __ldcg loads; the three buffers are reused every iteration
// without parity (each is written between one pair of barriers and read between the next).
#pragma once
#include <cooperative_groups.h>

#include "pcg_fused.cuh"

namespace gbdpcg {

template <typename T>
struct GridScratch {
    T *halo_r;     // [G][2][nx]  boundary block-rows of the initial residual (prologue)
    T *halo_rt;    // [G][2][nx]  boundary block-rows of r~
    T *halo_w;     // [G][2][nx]  boundary block-rows of w
    T *partials;   // [G][2]      per-CTA (r.r~, w.r~)
};
template <typename T>
__host__ __device__ inline size_t grid_scratch_elems(int nx, int G) { return (size_t)G * (6 * (size_t)nx + 2); }

struct GridPlan {
    uint32_t red_off, gsum_off, ghost_off, winA_off, winB_off, psm_off, stage_off, total;
};
template <typename T>
__host__ __device__ inline GridPlan make_grid_plan(int nx, int nxp, int bpc, int pmode)
{
    GridPlan p;
    uint32_t o = 0;
    p.red_off = o;   o += 2 * 2 * 32 * (uint32_t)sizeof(T);
    p.gsum_off = o;  o += 16 * 2;                                          // 2 slots x (gamma, delta)
    p.ghost_off = o; o += (uint32_t)round_up(4 * nx * (int)sizeof(T), 16);   // ghost r [2 sides][nx], ghost s [2 sides][nx]
    p.winA_off = o;  o += (uint32_t)round_up((bpc + 2) * nxp * (int)sizeof(T), 16);
    p.winB_off = o;  o += (uint32_t)round_up((bpc + 2) * nxp * (int)sizeof(T), 16);
    const uint32_t mat = (uint32_t)round_up(bpc * 3 * nx * nx * (int)sizeof(T), 16);
    p.psm_off = o;   o += (pmode == P_SMEM) ? mat : 0;
    p.stage_off = o; o += mat;                                             // one matrix at a time
    p.total = o;
    return p;
}

template <typename T, int NX, int R, int PMODE, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) pcg_grid_kernel(const KernelArgs<T> a, const GridScratch<T> gs)
{
    static_assert(NX % R == 0, "rows per thread must divide the state size");
    namespace cg = cooperative_groups;
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int BSZ = 3 * NX * NX;
    constexpr int TPB = NX / R;
    constexpr int NWP = round_up(round_up(MAXT, 32) / 32, 2);

    extern __shared__ __align__(16) unsigned char smem_raw[];
    cg::grid_group grid = cg::this_grid();
    const int N = (int)a.N, bpc = (int)a.bpc;
    const GridPlan plan = make_grid_plan<T>(NX, NXP, bpc, PMODE);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *gsum = reinterpret_cast<T *>(smem_raw + plan.gsum_off);
    T *ghost_r = reinterpret_cast<T *>(smem_raw + plan.ghost_off), *ghost_s = ghost_r + 2 * NX;
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *psm = reinterpret_cast<T *>(smem_raw + plan.psm_off);
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int cta = (int)blockIdx.x, G = (int)gridDim.x;
    const int lb = t / TPB, ti = t - lb * TPB, r0 = ti * R;
    const int b0 = cta * bpc;
    const int nb = max(0, min(bpc, N - b0));
    const bool active = lb < nb;
    const int b = b0 + lb;
    const bool has_left = cta > 0 && nb > 0, has_right = b0 + nb < N;
    // ghost threads: t < 2*NX handle element e of the ghost block-row on side `gside` (0 = left neighbour's last block-row,
    // 1 = right neighbour's first block-row)
    const int gside = t / NX, ge = t - gside * NX;
    const bool ghost_thread = t < 2 * NX && (gside == 0 ? has_left : has_right);
    const int gnb = gside == 0 ? cta - 1 : cta + 1;                       // neighbour CTA
    const size_t gsrc = ((size_t)gnb * 2 + (gside == 0 ? 1 : 0)) * NX + ge;   // its boundary block-row facing this CTA
    const int ghalo = (gside == 0 ? 0 : (nb + 1) * NXP) + ge;             // window halo slot

    for (int i = t; i < (bpc + 2) * NXP; i += blockDim.x) { winA[i] = (T)0; winB[i] = (T)0; }
    for (int i = t; i < 128; i += blockDim.x) red[i] = (T)0;
    __syncthreads();

    const int slot = (lb + 1) * NXP + r0;
    const int wofs = (active ? lb : 0) * NXP;
    const bool first_row = active && lb == 0, last_row = active && lb == nb - 1;
    auto publish = [&](T *win, const T (&x)[R]) {
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) win[slot + i] = x[i];
        }
    };
    // boundary block-rows of a vector -> global (side 0: first own block-row, side 1: last own block-row)
    auto send_boundary = [&](T *buf, const T (&x)[R]) {
        if (first_row) {
#pragma unroll
            for (int i = 0; i < R; i++) buf[((size_t)cta * 2 + 0) * NX + r0 + i] = x[i];
        }
        if (last_row) {
#pragma unroll
            for (int i = 0; i < R; i++) buf[((size_t)cta * 2 + 1) * NX + r0 + i] = x[i];
        }
    };
    int red_slot = 0;
    // CTA-wide sum of a pair, identical in every thread of the CTA
    auto cta_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 64;
        if (lane == 0) { rd[2 * warp] = p.x(); rd[2 * warp + 1] = p.y(); }
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, NWP>(rd);
        red_slot ^= 1;
        return s;
    };
    // grid-wide sum of the G per-CTA partials: warp 0 of EVERY CTA sums them in the same order (lane j takes j, j+32, ... in
    // sequence, then a butterfly), so the result is bit-identical in every CTA
    int gsum_slot = 0;
    auto grid_sum = [&]() -> Pair<T> {
        T *gd = gsum + gsum_slot * 2;
        if (warp == 0) {
            Pair<T> p((T)0, (T)0);
            for (int j = lane; j < G; j += 32) p = p + Pair<T>(__ldcg(gs.partials + 2 * j), __ldcg(gs.partials + 2 * j + 1));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
            if (lane == 0) { gd[0] = p.x(); gd[1] = p.y(); }
        }
        __syncthreads();
        Pair<T> s(gd[0], gd[1]);
        gsum_slot ^= 1;
        return s;
    };

    for (uint32_t prob = 0; prob < a.batch; prob++) {
        const size_t vbase = (size_t)prob * a.vec_stride;
        const size_t vofs = vbase + (size_t)b * NX + r0;
        T lam[R], gam[R];
#pragma unroll
        for (int i = 0; i < R; i++) { lam[i] = (T)0; gam[i] = (T)0; }
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) { lam[i] = a.lambda[vofs + i]; gam[i] = __ldg(a.gamma + vofs + i); }
        }

        // ---- matrices: HBM -> shared (cooperative coalesced loads) -> registers, one matrix at a time ----
        T Sreg[R][3 * NX];
        T Preg[(PMODE == P_REG) ? R : 1][(PMODE == P_REG) ? 3 * NX : 1];
        (void)Preg;
        auto load_rows = [&](T (&M)[R][3 * NX], const T *img) {
            const T *row = img + (active ? lb * BSZ + r0 : 0);
            const bool kill_left = !active || b == 0, kill_right = !active || b == N - 1, kill_mid = !active;
#pragma unroll
            for (int k = 0; k < 3 * NX; k++) {
                const bool kill = (k < NX) ? kill_left : (k >= 2 * NX) ? kill_right : kill_mid;   // padding blocks are never read (pcg.cuh:105-106)
#pragma unroll
                for (int i = 0; i < R; i++) M[i][k] = kill ? (T)0 : row[k * NX + i];
            }
        };
        stage_image<T>(a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ, stage, nb * BSZ);
        __syncthreads();
        load_rows(Sreg, stage);
        __syncthreads();
        if constexpr (PMODE == P_REG) {
            stage_image<T>(a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ, stage, nb * BSZ);
            __syncthreads();
            load_rows(Preg, stage);
            __syncthreads();
        } else if constexpr (PMODE == P_SMEM) {
            stage_image<T>(a.Pinv + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ, psm, nb * BSZ);
            __syncthreads();
            if (b0 == 0) for (int i = t; i < NX * NX; i += blockDim.x) psm[i] = (T)0;
            if (nb > 0 && b0 + nb == N) for (int i = t; i < NX * NX; i += blockDim.x) psm[(nb - 1) * BSZ + 2 * NX * NX + i] = (T)0;
            __syncthreads();
        }
        const T *Prow = psm + (active ? lb * BSZ + r0 : 0);
        (void)Prow;
        auto mv_S = [&](const T *win, T (&y)[R]) {
            block_row_mv<T, NX, R, false>(Sreg, nullptr, win + wofs, y);
            if (!active) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };
        auto mv_P = [&](const T *win, T (&y)[R]) {
            if constexpr (PMODE == P_REG) block_row_mv<T, NX, R, false>(Preg, nullptr, win + wofs, y);
            else if constexpr (PMODE == P_SMEM) block_row_mv<T, NX, R, true>(Sreg, Prow, win + wofs, y);
            else {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = win[wofs + NXP + r0 + i];       // identity: the middle block of the window
            }
            if (!active) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };

        // ---- r = gamma - S*lambda (pcg.cuh:118-126); the halo of lambda is read straight from the input vector ----
        T res[R], rt[R], w[R], p[R], s[R];
        publish(winB, lam);
        if (ghost_thread) winB[ghalo] = a.lambda[vbase + (size_t)(gside == 0 ? b0 - 1 : b0 + nb) * NX + ge];
        __syncthreads();
        mv_S(winB, w);
#pragma unroll
        for (int i = 0; i < R; i++) { res[i] = gam[i] - w[i]; p[i] = (T)0; s[i] = (T)0; }
        send_boundary(gs.halo_r, res);
        grid.sync();                                       // prologue barrier: initial ghost copies of r
        if (t < 2 * NX) {
            ghost_r[t] = ghost_thread ? __ldcg(gs.halo_r + gsrc) : (T)0;
            ghost_s[t] = (T)0;
        }

        uint32_t it = 0;
        bool hit_max = true;
        T rcp_gamma = (T)0, rcp_alpha = (T)0;
        for (;;) {
            publish(winA, res);
            if (ghost_thread) winA[ghalo] = ghost_r[t];
            __syncthreads();
            mv_P(winA, rt);                                // r~ = Phi^-1 r                (pcg.cuh:131-140, 181-184)
            publish(winB, rt);
            send_boundary(gs.halo_rt, rt);
            grid.sync();                                   // barrier 1: boundary rows of r~ are visible
            if (ghost_thread) winB[ghalo] = __ldcg(gs.halo_rt + gsrc);
            __syncthreads();
            mv_S(winB, w);                                 // w = S r~
            T g = (T)0, d = (T)0;
#pragma unroll
            for (int i = 0; i < R; i++) { g += res[i] * rt[i]; d += w[i] * rt[i]; }
            const Pair<T> cp = cta_sum(g, d);
            if (t == 0) { gs.partials[2 * cta] = cp.x(); gs.partials[2 * cta + 1] = cp.y(); }
            send_boundary(gs.halo_w, w);
            grid.sync();                                   // barrier 2: partials and boundary rows of w are visible
            const T wg = ghost_thread ? __ldcg(gs.halo_w + gsrc) : (T)0;   // (issued before the partials' round trip to L2, not after it)
            const Pair<T> gd = grid_sum();                 // gamma = r.r~ , delta = w.r~, bit-identical in every CTA
            const T gamma = gd.x(), delta = gd.y();
            if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
            if (it == a.max_iter) break;
            T beta, alpha;
            if (it == 0) { beta = (T)0; alpha = gamma / delta; }
            else {
                beta = gamma * rcp_gamma;
                alpha = gamma / (delta - beta * (gamma * rcp_alpha));
            }
#pragma unroll
            for (int i = 0; i < R; i++) {
                p[i] = fma_t(beta, p[i], rt[i]);
                lam[i] = fma_t(alpha, p[i], lam[i]);
                s[i] = fma_t(beta, s[i], w[i]);
                res[i] = fma_t(-alpha, s[i], res[i]);
            }
            if (ghost_thread) {                            // the same two fused multiply-adds on the ghost rows
                const T sg = fma_t(beta, ghost_s[t], wg);
                ghost_s[t] = sg;
                ghost_r[t] = fma_t(-alpha, sg, ghost_r[t]);
            }
            rcp_gamma = (T)1 / gamma;
            rcp_alpha = (T)1 / alpha;
            ++it;
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) a.lambda[vofs + i] = lam[i];
        }
        if (cta == 0 && t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        __syncthreads();                                   // staging, windows and ghosts are reused by the next system
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// pcg_grid1_kernel: the same solve with ONE grid barrier per iteration.
//
// The first barrier of pcg_grid_kernel exists only to hand the neighbours' boundary block-row of r~ = Phi^-1 r across.  Here a CTA
// computes that row itself: it keeps the neighbour's boundary block-row of Phi^-1 in the registers of TPB otherwise idle threads
// ("ghost rows": rows of Phi^-1 only, S rows zero, so they add nothing to w or to the dot products) and ghost copies of r and s
// TWO block-rows deep on either side, so the window of a ghost row -- r on block-rows b0-2, b0-1, b0 -- is always at hand.  The
// ghost row runs through the very same block_row_mv as on its owner, so the local r~ is bit-identical to what the exchange
// delivered.  What still crosses CTAs, with the one remaining barrier: the per-CTA partials and the boundary rows of w, now two
// block-rows per side (the second level keeps the second-level ghosts of s and r current).  Because a fast CTA may be one
// iteration ahead of a slow one between two barriers, the exchange buffers alternate with the iteration's parity.
// Needs bpc >= 2 (a neighbour must own both block-rows a side looks at) and 2 TPB threads beyond the owners' (planner).
// Scratch (elements): hw [2 parities][G][2 sides][2 levels][nx] | partials [2 parities][G][2].
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__host__ __device__ inline size_t grid1_scratch_elems(int nx, int G) { return (size_t)G * (8 * (size_t)nx + 4); }

template <typename T>
__host__ __device__ inline GridPlan make_grid1_plan(int nx, int nxp, int bpc, int pmode)
{
    GridPlan p;
    uint32_t o = 0;
    p.red_off = o;   o += 2 * 2 * 32 * (uint32_t)sizeof(T);
    p.gsum_off = o;  o += 16 * 2;
    p.ghost_off = o; o += (uint32_t)round_up(8 * nx * (int)sizeof(T), 16);   // ghost r, ghost s: [2 sides][2 levels][nx] each
    p.winA_off = o;  o += (uint32_t)round_up((bpc + 4) * nxp * (int)sizeof(T), 16);
    p.winB_off = o;  o += (uint32_t)round_up((bpc + 4) * nxp * (int)sizeof(T), 16);
    const uint32_t mat = (uint32_t)round_up((bpc + 2) * 3 * nx * nx * (int)sizeof(T), 16);   // own block-rows + one ghost row per side
    p.psm_off = o;   o += (pmode == P_SMEM) ? mat : 0;
    p.stage_off = o; o += mat;
    p.total = o;
    return p;
}

template <typename T, int NX, int R, int PMODE, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) pcg_grid1_kernel(const KernelArgs<T> a, T *const scratch)
{
    static_assert(NX % R == 0, "rows per thread must divide the state size");
    namespace cg = cooperative_groups;
    constexpr int NXP = Geo<T>::nxp(NX);
    constexpr int BSZ = 3 * NX * NX;
    constexpr int TPB = NX / R;
    constexpr int NWP = round_up(round_up(MAXT, 32) / 32, 2);

    extern __shared__ __align__(16) unsigned char smem_raw[];
    cg::grid_group grid = cg::this_grid();
    const int N = (int)a.N, bpc = (int)a.bpc;
    const GridPlan plan = make_grid1_plan<T>(NX, NXP, bpc, PMODE);
    T *red = reinterpret_cast<T *>(smem_raw + plan.red_off);
    T *gsum = reinterpret_cast<T *>(smem_raw + plan.gsum_off);
    T *ghost_r = reinterpret_cast<T *>(smem_raw + plan.ghost_off), *ghost_s = ghost_r + 4 * NX;
    T *winA = reinterpret_cast<T *>(smem_raw + plan.winA_off);
    T *winB = reinterpret_cast<T *>(smem_raw + plan.winB_off);
    T *psm = reinterpret_cast<T *>(smem_raw + plan.psm_off);
    T *stage = reinterpret_cast<T *>(smem_raw + plan.stage_off);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int cta = (int)blockIdx.x, G = (int)gridDim.x;
    const int lb = t / TPB, ti = t - lb * TPB, r0 = ti * R;
    const int b0 = cta * bpc;
    const int nb = max(0, min(bpc, N - b0));
    const bool active = lb < nb;
    const int b = b0 + lb;
    const bool has_left = cta > 0 && nb > 0, has_right = b0 + nb < N, has_right2 = b0 + nb + 1 < N;
    // ghost ROWS: the TPB threads after the owners' take the left neighbour's last block-row of Phi^-1, the next TPB the right
    // neighbour's first one (bpc * TPB is a multiple of TPB, so such a thread's row offset is r0 as well)
    const int gt = t - bpc * TPB;
    const int gmv_side = (gt >= TPB) ? 1 : 0;
    const bool gmv = gt >= 0 && gt < 2 * TPB && (gmv_side == 0 ? has_left : has_right);
    const int grow = gmv_side == 0 ? b0 - 1 : b0 + nb;                      // the ghost row's block-row in the system
    const bool prow_mine = active || gmv;                                  // this thread owns rows of Phi^-1
    // ghost ELEMENTS: thread t < 4 NX keeps element ge of (side, level): level 0 = the adjacent block-row, 1 = the one behind it
    const int gq = t / NX, ge = t - gq * NX, gside = (gq >> 1) & 1, glvl = gq & 1;
    const bool gel = t < 4 * NX && (gside == 0 ? has_left : (glvl == 0 ? has_right : has_right2));
    const int gnb = gside == 0 ? cta - 1 : cta + 1;                        // neighbour CTA; its side facing this CTA: 1 (last rows) / 0 (first rows)
    const size_t gsrc = ((((size_t)gnb * 2 + (gside == 0 ? 1 : 0)) * 2) + (size_t)glvl) * NX + ge;
    const int gslot = (gside == 0 ? (1 - glvl) : (nb + 2 + glvl)) * NXP + ge;   // window: [L2][L1][own ...][R1][R2]
    const size_t hw_par = (size_t)G * 4 * NX;                              // elements per parity of hw
    T *const hw = scratch;
    T *const partials = scratch + 2 * hw_par;

    for (int i = t; i < (bpc + 4) * NXP; i += blockDim.x) { winA[i] = (T)0; winB[i] = (T)0; }
    for (int i = t; i < 128; i += blockDim.x) red[i] = (T)0;
    __syncthreads();

    const int slot = (lb + 2) * NXP + r0;                                  // own rows
    const int gmv_slot = (gmv_side == 0 ? 1 : nb + 2) * NXP + r0;          // ghost row's r~
    const int wofs = (active ? lb + 1 : 0) * NXP;
    const int pwofs = active ? wofs : (gmv ? (gmv_side == 0 ? 0 : (nb + 1) * NXP) : 0);
    auto publish = [&](T *win, const T (&x)[R]) {
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) win[slot + i] = x[i];
        }
    };
    // the two boundary block-rows of either side -> global: (side 0: first / second own block-row, side 1: last / last but one)
    auto send_boundary = [&](T *buf, const T (&x)[R]) {
        if (active) {
            const int lv0 = lb, lv1 = nb - 1 - lb;                         // level when seen from the left / from the right
            if (lv0 < 2) {
#pragma unroll
                for (int i = 0; i < R; i++) buf[(((size_t)cta * 2 + 0) * 2 + lv0) * NX + r0 + i] = x[i];
            }
            if (lv1 < 2) {
#pragma unroll
                for (int i = 0; i < R; i++) buf[(((size_t)cta * 2 + 1) * 2 + lv1) * NX + r0 + i] = x[i];
            }
        }
    };
    int red_slot = 0;
    auto cta_sum = [&](T x0, T x1) -> Pair<T> {
        Pair<T> p(x0, x1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
        T *rd = red + red_slot * 64;
        if (lane == 0) { rd[2 * warp] = p.x(); rd[2 * warp + 1] = p.y(); }
        __syncthreads();
        Pair<T> s = pair_tree_sum<T, NWP>(rd);
        red_slot ^= 1;
        return s;
    };
    int gsum_slot = 0;
    auto grid_sum = [&](const T *part) -> Pair<T> {
        T *gd = gsum + gsum_slot * 2;
        if (warp == 0) {
            Pair<T> p((T)0, (T)0);
            for (int j = lane; j < G; j += 32) p = p + Pair<T>(__ldcg(part + 2 * j), __ldcg(part + 2 * j + 1));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) p = p + shfl_xor_pair<T>(p, o);
            if (lane == 0) { gd[0] = p.x(); gd[1] = p.y(); }
        }
        __syncthreads();
        Pair<T> s(gd[0], gd[1]);
        gsum_slot ^= 1;
        return s;
    };

    for (uint32_t prob = 0; prob < a.batch; prob++) {
        const size_t vbase = (size_t)prob * a.vec_stride;
        const size_t vofs = vbase + (size_t)b * NX + r0;
        T lam[R], gam[R];
#pragma unroll
        for (int i = 0; i < R; i++) { lam[i] = (T)0; gam[i] = (T)0; }
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) { lam[i] = a.lambda[vofs + i]; gam[i] = __ldg(a.gamma + vofs + i); }
        }

        // ---- matrices: HBM -> shared -> registers, one matrix at a time.  Image row i of the staging buffer is block-row
        //      b0 - 1 + i: the owners read rows 1..nb, the ghost rows of Phi^-1 rows 0 and nb + 1 ----
        T Sreg[R][3 * NX];
        T Preg[(PMODE == P_REG) ? R : 1][(PMODE == P_REG) ? 3 * NX : 1];
        (void)Preg;
        const int irow = active ? lb + 1 : (gmv_side == 0 ? 0 : nb + 1);
        auto load_rows = [&](T (&M)[R][3 * NX], const T *img, const bool with_ghost) {
            const bool mine = active || (with_ghost && gmv);
            const int brow = active ? b : grow;
            const T *row = img + (mine ? irow * BSZ + r0 : 0);
            const bool kill_left = !mine || brow == 0, kill_right = !mine || brow == N - 1, kill_mid = !mine;
#pragma unroll
            for (int k = 0; k < 3 * NX; k++) {
                const bool kill = (k < NX) ? kill_left : (k >= 2 * NX) ? kill_right : kill_mid;   // padding blocks are never read (pcg.cuh:105-106)
#pragma unroll
                for (int i = 0; i < R; i++) M[i][k] = kill ? (T)0 : row[k * NX + i];
            }
        };
        stage_image<T>(a.S + (size_t)prob * a.mat_stride + (size_t)b0 * BSZ, stage + BSZ, nb * BSZ);
        __syncthreads();
        load_rows(Sreg, stage, false);
        __syncthreads();
        const int gl = has_left ? 1 : 0, gr = has_right ? 1 : 0;           // ghost rows that exist
        if constexpr (PMODE == P_REG) {
            stage_image<T>(a.Pinv + (size_t)prob * a.mat_stride + (size_t)(b0 - gl) * BSZ, stage + (1 - gl) * BSZ, (nb + gl + gr) * BSZ);
            __syncthreads();
            load_rows(Preg, stage, true);
            __syncthreads();
        } else if constexpr (PMODE == P_SMEM) {
            stage_image<T>(a.Pinv + (size_t)prob * a.mat_stride + (size_t)(b0 - gl) * BSZ, psm + (1 - gl) * BSZ, (nb + gl + gr) * BSZ);
            __syncthreads();
            // padding blocks: left block of block-row 0 (image row 1 of CTA 0), right block of block-row N-1 (an own row or the right ghost row)
            if (b0 == 0) for (int i = t; i < NX * NX; i += blockDim.x) psm[BSZ + i] = (T)0;
            if (nb > 0 && b0 + nb == N) for (int i = t; i < NX * NX; i += blockDim.x) psm[nb * BSZ + 2 * NX * NX + i] = (T)0;
            if (b0 + nb == N - 1) for (int i = t; i < NX * NX; i += blockDim.x) psm[(nb + 1) * BSZ + 2 * NX * NX + i] = (T)0;
            __syncthreads();
        }
        const T *Prow = psm + (prow_mine ? irow * BSZ + r0 : 0);
        (void)Prow;
        auto mv_S = [&](const T *win, T (&y)[R]) {
            block_row_mv<T, NX, R, false>(Sreg, nullptr, win + wofs, y);
            if (!active) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };
        // r~ rows: the owners' and, with the same code on the same operands as on their owner, the two ghost rows
        auto mv_P = [&](const T *win, T (&y)[R]) {
            if constexpr (PMODE == P_REG) block_row_mv<T, NX, R, false>(Preg, nullptr, win + pwofs, y);
            else if constexpr (PMODE == P_SMEM) block_row_mv<T, NX, R, true>(Sreg, Prow, win + pwofs, y);
            else {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = win[pwofs + NXP + r0 + i];      // identity: the middle block of the window
            }
            if (!prow_mine) {
#pragma unroll
                for (int i = 0; i < R; i++) y[i] = (T)0;
            }
        };

        // ---- r = gamma - S*lambda (pcg.cuh:118-126); the halo of lambda is read straight from the input vector ----
        T res[R], rt[R], w[R], p[R], s[R];
        publish(winB, lam);
        if (gel && glvl == 0) winB[gslot] = a.lambda[vbase + (size_t)(gside == 0 ? b0 - 1 : b0 + nb) * NX + ge];
        __syncthreads();
        mv_S(winB, w);
#pragma unroll
        for (int i = 0; i < R; i++) { res[i] = gam[i] - w[i]; p[i] = (T)0; s[i] = (T)0; }
        send_boundary(hw + hw_par, res);                   // (parity 1: iteration 0 writes parity 0)
        grid.sync();                                       // prologue barrier: initial ghost copies of r, two block-rows deep
        if (t < 4 * NX) {
            ghost_r[t] = gel ? __ldcg(hw + hw_par + gsrc) : (T)0;
            ghost_s[t] = (T)0;
        }

        uint32_t it = 0, par = 0;
        bool hit_max = true;
        T rcp_gamma = (T)0, rcp_alpha = (T)0;
        for (;;) {
            publish(winA, res);
            if (gel) winA[gslot] = ghost_r[t];
            __syncthreads();
            mv_P(winA, rt);                                // r~ = Phi^-1 r, ghost rows included     (pcg.cuh:131-140, 181-184)
            publish(winB, rt);
            if (gmv) {
#pragma unroll
                for (int i = 0; i < R; i++) winB[gmv_slot + i] = rt[i];
            }
            __syncthreads();
            mv_S(winB, w);                                 // w = S r~
            T *const part = partials + (size_t)par * 2 * G;
            T *const hwp = hw + (size_t)par * hw_par;
            send_boundary(hwp, w);                         // (on their way while the CTA reduces)
            T g = (T)0, d = (T)0;
            if (active) {                                  // (a ghost row's r~ is not part of this CTA's share of the dot products)
#pragma unroll
                for (int i = 0; i < R; i++) { g += res[i] * rt[i]; d += w[i] * rt[i]; }
            }
            const Pair<T> cp = cta_sum(g, d);
            if (t == 0) { part[2 * cta] = cp.x(); part[2 * cta + 1] = cp.y(); }
            grid.sync();                                   // THE barrier: partials and boundary rows of w are visible
            const T wg = gel ? __ldcg(hwp + gsrc) : (T)0;  // (issued before the partials' round trip to L2, not after it)
            const Pair<T> gd = grid_sum(part);             // gamma = r.r~ , delta = w.r~, bit-identical in every CTA
            const T gamma = gd.x(), delta = gd.y();
            if (it > 0 && abs_t(gamma) < a.tol) { hit_max = false; break; }   // reference exit test (pcg.cuh:195)
            if (it == a.max_iter) break;
            T beta, alpha;
            if (it == 0) { beta = (T)0; alpha = gamma / delta; }
            else {
                beta = gamma * rcp_gamma;
                alpha = gamma / (delta - beta * (gamma * rcp_alpha));
            }
            if (active) {
#pragma unroll
                for (int i = 0; i < R; i++) {
                    p[i] = fma_t(beta, p[i], rt[i]);
                    lam[i] = fma_t(alpha, p[i], lam[i]);
                    s[i] = fma_t(beta, s[i], w[i]);
                    res[i] = fma_t(-alpha, s[i], res[i]);
                }
            }
            if (gel) {                                     // the same two fused multiply-adds on the ghost elements, both levels
                const T sg = fma_t(beta, ghost_s[t], wg);
                ghost_s[t] = sg;
                ghost_r[t] = fma_t(-alpha, sg, ghost_r[t]);
            }
            rcp_gamma = (T)1 / gamma;
            rcp_alpha = (T)1 / alpha;
            par ^= 1u;
            ++it;
        }

        // ---- outputs                                                            (pcg.cuh:212-215) ----
        if (active) {
#pragma unroll
            for (int i = 0; i < R; i++) a.lambda[vofs + i] = lam[i];
        }
        if (cta == 0 && t == 0) {
            if (a.iters) a.iters[prob] = it;
            if (a.exits) a.exits[prob] = hit_max ? 1 : 0;
        }
        __syncthreads();                                   // staging, windows and ghosts are reused by the next system
    }
}

}  // namespace gbdpcg
