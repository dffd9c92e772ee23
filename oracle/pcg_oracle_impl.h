/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference block-tridiagonal PCG.
 *
 * This file is included twice by pcg_oracle.c, once with T=float (SUF=f32) and once
 * with T=double (SUF=f64).  It is never linked into the product library.
 *
 * It restates, on the host and for one problem at a time:
 *   - the compressed block-tridiagonal block-row mat-vec     reference include/utils.cuh:46-85  (bdmv)
 *   - the 3*nx halo window of a vector                       reference include/utils.cuh:9-40   (loadbdVec)
 *   - the PCG recurrence, exit test and outputs              reference include/pcg.cuh:118-212
 *   - (order==1 only) the shared-memory tree reduction order reference GLASS/src/L1/reduce.cuh:35-62, 64-96
 *                                                            and the dot of GLASS/src/L1/dot.cuh:49-60
 *
 * Parity pinning: the reference ships no golden vectors or tests; this
 * restatement is pinned against (a) the hand-derived C1 known answers (tests/golden/c1_kat.json)
 * and (b) outputs of the reference CUDA kernel itself, compiled unmodified from
 * /root/reference by oracle/Makefile into oracle/_ref/ and run on a B200
 * (tests/golden/ref_cuda_*.npz, generator oracle/make_ref_fixtures.py).
 */

#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

/* y_b = M_b . [x_{b-1} | x_b | x_{b+1}], accumulating columns in ascending order like
 * bdmv (utils.cuh:58-84): block 0 skips its left block, block N-1 skips its right block,
 * so the padding blocks of the layout are never touched. */
static void FN(blockrow_mv)(const T *M, const T *x, T *y, uint32_t nx, uint32_t N, uint32_t b)
{
    const T *Mb = M + (size_t)b * 3u * nx * nx;
    uint32_t c_lo = (b == 0) ? nx : 0;
    uint32_t c_hi = (b == N - 1) ? 2u * nx : 3u * nx;
    for (uint32_t r = 0; r < nx; r++) {
        T acc = (T)0;
        for (uint32_t c = c_lo; c < c_hi; c++) {
            /* window column c maps to global vector element (b-1)*nx + c */
            acc += Mb[(size_t)nx * c + r] * x[(size_t)b * nx + c - nx];
        }
        y[(size_t)b * nx + r] = acc;
    }
}

void FN(oracle_bdmv)(const T *M, const T *x, T *y, uint32_t nx, uint32_t N)
{
    for (uint32_t b = 0; b < N; b++) FN(blockrow_mv)(M, x, y, nx, N, b);
}

/* GLASS tree reduction, in place (reduce.cuh:35-62).  Sequential emulation: within one
 * halving step the reads x[i+s] are of elements no thread writes in that step, so a plain
 * ascending loop reproduces the parallel result; the odd tail is folded into x[0] after
 * x[0]'s own add (same thread, program order). */
static T FN(glass_reduce)(T *x, uint32_t n)
{
    uint32_t s = n;
    while (s > 3) {
        uint32_t odd = s & 1u;
        s = (s - odd) / 2u;
        for (uint32_t i = 0; i < s; i++) x[i] += x[i + s];
        if (odd) x[0] += x[2u * s];
    }
    for (uint32_t i = 1; i < s; i++) x[0] += x[i];
    return x[0];
}

/* Global dot product a.b over N blocks of nx.
 * order 0: plain left-to-right accumulation (the "straightforward host oracle").
 * order 1: reference order -- per block dot via glass::dot (products then tree), then
 *          the N per-block partials through the same tree (pcg.cuh:144-148,163-168,187-192). */
static T FN(dot)(const T *a, const T *b, uint32_t nx, uint32_t N, int order, T *scratch)
{
    if (order == 0) {
        T acc = (T)0;
        for (size_t i = 0; i < (size_t)nx * N; i++) acc += a[i] * b[i];
        return acc;
    }
    T *part = scratch;            /* N entries  */
    T *blk = scratch + N;         /* nx entries */
    for (uint32_t k = 0; k < N; k++) {
        for (uint32_t i = 0; i < nx; i++) blk[i] = a[(size_t)k * nx + i] * b[(size_t)k * nx + i];
        part[k] = FN(glass_reduce)(blk, nx);
    }
    return FN(glass_reduce)(part, N);
}

/*
 * One PCG solve.  Pinv == NULL means identity preconditioner (r_tilde = r).
 * lambda is in/out (warm start).  Returns 0, or -1 on bad arguments / allocation failure.
 * eta_trace (nullable, capacity max_iter+1) receives eta_0, eta'_1, ... for diagnostics.
 * Follows pcg.cuh:118-212 line by line:
 *   r = gamma - S lambda ; rt = Pinv r ; p = rt ; eta = r.rt
 *   loop: ups = S p ; alpha = eta/(p.ups) ; lambda += alpha p ; r -= alpha ups ;
 *         rt = Pinv r ; eta' = r.rt ; if |eta'| < tol {iter++; exit=false; break}
 *         beta = eta'/eta ; eta = eta' ; p = rt + beta p
 */
int FN(oracle_pcg)(const T *S, const T *Pinv, const T *gamma, T *lambda,
                   uint32_t nx, uint32_t N, T exit_tol, uint32_t max_iter, int order,
                   uint32_t *iters_out, int *max_iter_exit_out, T *eta_trace)
{
    if (!S || !gamma || !lambda || nx == 0 || N < 2) return -1;
    size_t n = (size_t)nx * N;
    T *buf = (T *)malloc(sizeof(T) * (4 * n + N + nx));
    if (!buf) return -1;
    T *r = buf, *rt = buf + n, *p = buf + 2 * n, *ups = buf + 3 * n, *scratch = buf + 4 * n;

    FN(oracle_bdmv)(S, lambda, r, nx, N);
    for (size_t i = 0; i < n; i++) r[i] = gamma[i] - r[i];
    if (Pinv) FN(oracle_bdmv)(Pinv, r, rt, nx, N);
    else memcpy(rt, r, sizeof(T) * n);
    memcpy(p, rt, sizeof(T) * n);
    T eta = FN(dot)(r, rt, nx, N, order, scratch);
    if (eta_trace) eta_trace[0] = eta;

    uint32_t iter;
    int max_iter_exit = 1;
    for (iter = 0; iter < max_iter; iter++) {
        FN(oracle_bdmv)(S, p, ups, nx, N);
        T v = FN(dot)(p, ups, nx, N, order, scratch);
        T alpha = eta / v;
        for (size_t i = 0; i < n; i++) {
            lambda[i] += alpha * p[i];
            r[i] -= alpha * ups[i];
        }
        if (Pinv) FN(oracle_bdmv)(Pinv, r, rt, nx, N);
        else memcpy(rt, r, sizeof(T) * n);
        T eta_new = FN(dot)(r, rt, nx, N, order, scratch);
        if (eta_trace) eta_trace[iter + 1] = eta_new;
        if (FABS(eta_new) < exit_tol) { iter++; max_iter_exit = 0; break; }
        T beta = eta_new / eta;
        eta = eta_new;
        for (size_t i = 0; i < n; i++) p[i] = rt[i] + beta * p[i];
    }
    if (iters_out) *iters_out = iter;
    if (max_iter_exit_out) *max_iter_exit_out = max_iter_exit;
    free(buf);
    return 0;
}

/* ---- batched driver over a pthread pool: problem i uses slabs at i*3*N*nx^2 / i*N*nx ---- */
typedef struct {
    const T *S, *Pinv, *gamma;
    T *lambda;
    uint32_t nx, N, batch, max_iter;
    T tol;
    int order;
    uint32_t *iters;
    uint8_t *exits;
    volatile uint32_t *next;
} FN(batch_job);

static void *FN(batch_worker)(void *arg)
{
    FN(batch_job) *j = (FN(batch_job) *)arg;
    size_t ms = (size_t)3 * j->N * j->nx * j->nx, vs = (size_t)j->N * j->nx;
    for (;;) {
        uint32_t i = __atomic_fetch_add(j->next, 1u, __ATOMIC_RELAXED);
        if (i >= j->batch) break;
        uint32_t it = 0;
        int ex = 0;
        FN(oracle_pcg)(j->S + i * ms, j->Pinv ? j->Pinv + i * ms : NULL, j->gamma + i * vs,
                       j->lambda + i * vs, j->nx, j->N, j->tol, j->max_iter, j->order, &it, &ex, NULL);
        if (j->iters) j->iters[i] = it;
        if (j->exits) j->exits[i] = (uint8_t)ex;
    }
    return NULL;
}

int FN(oracle_pcg_batched)(const T *S, const T *Pinv, const T *gamma, T *lambda,
                           uint32_t nx, uint32_t N, uint32_t batch, T exit_tol, uint32_t max_iter,
                           int order, uint32_t *iters, uint8_t *exits, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    volatile uint32_t next = 0;
    FN(batch_job) job = {S, Pinv, gamma, lambda, nx, N, batch, max_iter, exit_tol, order, iters, exits, &next};
    pthread_t th[256];
    int started = 0;
    for (int t = 1; t < threads; t++)
        if (pthread_create(&th[started], NULL, FN(batch_worker), &job) == 0) started++;
    FN(batch_worker)(&job);
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Preconditioner construction.  The reference never calls this, but it
 * carries the primitive: glass::invertMatrix(dimA, A, s_temp), GLASS/src/L3/inv.cuh:6-37 --
 * Gauss-Jordan on the augmented [V | I] (column-major, nx x 2nx) WITHOUT pivoting; per pivot p
 * the window of columns p .. p+nx is updated: pivot row *= pvInv, every other row
 * orv -= orpcv * pvInv * prvOld (left to right).  Restated literally; the inverse is the right
 * half.  Formulae for the blocks: the paper cited at reference README.md:66-76 (not in the tree).
 *   kind 0: Phi^-1 = I;  1: Phi^-1_kk = S_kk^-1;
 *   2: additionally Phi^-1_{k,k+-1} = -(S_kk^-1 (S_{k,k+-1} S_{k+-1,k+-1}^-1))   (symmetric stair)
 * Parity unpinned by the reference (no caller, no vectors): pinned by dense-inverse properties
 * in tests/test_pinv_cpu.py (P_kk S_kk = I; stair blocks against numpy.linalg.inv in fp64).
 * --------------------------------------------------------------------------------------------- */
static void FN(invert_block)(const T *A, T *Ainv, uint32_t n, T *aug, T *s_temp)
{
    /* aug = [A | I], column-major, n rows x 2n columns */
    for (uint32_t c = 0; c < n; c++)
        for (uint32_t r = 0; r < n; r++) {
            aug[(size_t)c * n + r] = A[(size_t)c * n + r];
            aug[(size_t)(n + c) * n + r] = (r == c) ? (T)1 : (T)0;
        }
    for (uint32_t piv = 0; piv < n; piv++) {
        size_t pco = (size_t)piv * n;
        T pvInv = (T)1 / aug[piv + pco];
        for (uint32_t ind = 0; ind < 2 * n + 1; ind++)               /* inv.cuh:18-23: pivot column, then pivot row */
            s_temp[ind] = (ind < n) ? aug[ind + pco] : aug[piv + pco + (size_t)(ind - n) * n];
        for (uint32_t ind = 0; ind < n * (n + 1); ind++) {           /* inv.cuh:26-31 */
            uint32_t row = ind % n, col = ind / n;
            size_t at = row + pco + (size_t)col * n;
            if (row == piv) aug[at] *= pvInv;
            else aug[at] -= s_temp[row] * pvInv * s_temp[n + col];
        }
    }
    for (uint32_t e = 0; e < n * n; e++) Ainv[e] = aug[(size_t)n * n + e];
}

/* C = A B, all n x n column-major, inner index ascending */
static void FN(block_mm)(const T *A, const T *B, T *Cm, uint32_t n)
{
    for (uint32_t c = 0; c < n; c++)
        for (uint32_t r = 0; r < n; r++) {
            T acc = (T)0;
            for (uint32_t q = 0; q < n; q++) acc += A[(size_t)q * n + r] * B[(size_t)c * n + q];
            Cm[(size_t)c * n + r] = acc;
        }
}

int FN(oracle_build_pinv)(const T *S, T *P, uint32_t nx, uint32_t N, int kind)
{
    if (!S || !P || nx == 0 || N < 2 || kind < 0 || kind > 2) return -1;
    size_t bb = (size_t)nx * nx, bsz = 3 * bb;
    T *dinv = (T *)malloc(sizeof(T) * bb * N);
    T *aug = (T *)malloc(sizeof(T) * 2 * bb + sizeof(T) * nx);
    T *tmp = (T *)malloc(sizeof(T) * 2 * bb);
    T *s_temp = (T *)malloc(sizeof(T) * (2 * nx + 1));
    if (!dinv || !aug || !tmp || !s_temp) { free(dinv); free(aug); free(tmp); free(s_temp); return -2; }
    memset(P, 0, sizeof(T) * bsz * N);
    for (uint32_t k = 0; k < N; k++) {
        if (kind == 0) {
            for (uint32_t r = 0; r < nx; r++) P[k * bsz + bb + (size_t)r * nx + r] = (T)1;
        } else {
            FN(invert_block)(S + k * bsz + bb, dinv + k * bb, nx, aug, s_temp);
            memcpy(P + k * bsz + bb, dinv + k * bb, sizeof(T) * bb);
        }
    }
    if (kind == 2) {
        for (uint32_t k = 0; k < N; k++)
            for (int side = 0; side < 2; side++) {
                if ((side == 0 && k == 0) || (side == 1 && k == N - 1)) continue;
                uint32_t kn = side ? k + 1 : k - 1;
                const T *So = S + k * bsz + (side ? 2 * bb : 0);
                T *dst = P + k * bsz + (side ? 2 * bb : 0);
                FN(block_mm)(So, dinv + kn * bb, tmp, nx);
                FN(block_mm)(dinv + k * bb, tmp, tmp + bb, nx);
                for (size_t e = 0; e < bb; e++) dst[e] = -tmp[bb + e];
            }
    }
    free(dinv); free(aug); free(tmp); free(s_temp);
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Schur-complement formation and primal recovery (the steps either side
 * of the solve).  Not in the reference tree (upstream MPCGPU, reference README.md:66-76); the
 * construction is the one the reference example decodes to and that
 * gbd_pcg_b200/synth.py restates on the host.  Formulae and layouts: include/gbdpcg.h.
 * Same arithmetic order as the CUDA kernel (gbd_pcg_b200/csrc/schur_form.cuh): Gauss-Jordan in
 * glass::invertMatrix order, inner indices ascending, diag = -((AQ A^T + BR B^T) + Qj^-1),
 * gamma_j = ((Qj^-1 q_j - AQ q_k) - BR r_k) - c_j.
 * Parity unpinned by the reference; pinned against a dense KKT solve in tests/test_schur_cpu.py.
 * --------------------------------------------------------------------------------------------- */
int FN(oracle_schur_form)(const T *Q, const T *R, const T *A, const T *B, const T *q, const T *r, const T *c,
                          uint32_t nx, uint32_t nu, uint32_t N, T *S, T *gamma, T *Qinv, T *Rinv)
{
    if (!Q || !R || !A || !B || !q || !r || !c || !S || !gamma || nx == 0 || nu == 0 || N < 2) return -1;
    size_t bb = (size_t)nx * nx, uu = (size_t)nu * nu, xu = (size_t)nx * nu, m = nx > nu ? nx : nu;
    T *Qi = (T *)malloc(sizeof(T) * bb * N), *Ri = (T *)malloc(sizeof(T) * uu * (N - 1));
    T *aug = (T *)malloc(sizeof(T) * (2 * m * m + m)), *s_temp = (T *)malloc(sizeof(T) * (2 * m + 1));
    T *AQ = (T *)malloc(sizeof(T) * bb * (N - 1)), *BR = (T *)malloc(sizeof(T) * xu);
    if (!Qi || !Ri || !aug || !s_temp || !AQ || !BR) { free(Qi); free(Ri); free(aug); free(s_temp); free(AQ); free(BR); return -2; }
    for (uint32_t k = 0; k < N; k++) FN(invert_block)(Q + k * bb, Qi + k * bb, nx, aug, s_temp);
    for (uint32_t k = 0; k + 1 < N; k++) FN(invert_block)(R + k * uu, Ri + k * uu, nu, aug, s_temp);
    for (uint32_t k = 0; k + 1 < N; k++) FN(block_mm)(A + k * bb, Qi + k * bb, AQ + k * bb, nx);      /* A_k Q_k^-1 */
    memset(S, 0, sizeof(T) * 3 * bb * N);
    for (uint32_t j = 0; j < N; j++) {
        T *Sr = S + (size_t)j * 3 * bb;
        const T *Qj = Qi + j * bb;
        if (j + 1 < N)                                                                              /* right = (A_j Q_j^-1)^T */
            for (uint32_t cc = 0; cc < nx; cc++)
                for (uint32_t rr = 0; rr < nx; rr++) Sr[2 * bb + (size_t)cc * nx + rr] = AQ[j * bb + (size_t)rr * nx + cc];
        if (j == 0) {
            for (size_t e = 0; e < bb; e++) Sr[bb + e] = -Qj[e];
            for (uint32_t e = 0; e < nx; e++) {
                T s = (T)0;
                for (uint32_t p = 0; p < nx; p++) s += Qj[(size_t)p * nx + e] * q[p];
                gamma[e] = s - c[e];
            }
            continue;
        }
        uint32_t k = j - 1;
        const T *AQk = AQ + k * bb, *Ak = A + k * bb, *Bk = B + k * xu, *Rk = Ri + k * uu;
        for (uint32_t cc = 0; cc < nu; cc++)                                                        /* BR = B_k R_k^-1 */
            for (uint32_t rr = 0; rr < nx; rr++) {
                T s = (T)0;
                for (uint32_t p = 0; p < nu; p++) s += Bk[(size_t)p * nx + rr] * Rk[(size_t)cc * nu + p];
                BR[(size_t)cc * nx + rr] = s;
            }
        memcpy(Sr, AQk, sizeof(T) * bb);                                                            /* left */
        for (uint32_t cc = 0; cc < nx; cc++)
            for (uint32_t rr = 0; rr < nx; rr++) {
                T s = (T)0, s2 = (T)0;
                for (uint32_t p = 0; p < nx; p++) s += AQk[(size_t)p * nx + rr] * Ak[(size_t)p * nx + cc];
                for (uint32_t p = 0; p < nu; p++) s2 += BR[(size_t)p * nx + rr] * Bk[(size_t)p * nx + cc];
                Sr[bb + (size_t)cc * nx + rr] = -((s + s2) + Qj[(size_t)cc * nx + rr]);
            }
        for (uint32_t e = 0; e < nx; e++) {
            T s = (T)0, s1 = (T)0, s2 = (T)0;
            for (uint32_t p = 0; p < nx; p++) { s += Qj[(size_t)p * nx + e] * q[(size_t)j * nx + p]; s1 += AQk[(size_t)p * nx + e] * q[(size_t)k * nx + p]; }
            for (uint32_t p = 0; p < nu; p++) s2 += BR[(size_t)p * nx + e] * r[(size_t)k * nu + p];
            gamma[(size_t)j * nx + e] = ((s - s1) - s2) - c[(size_t)j * nx + e];
        }
    }
    if (Qinv) memcpy(Qinv, Qi, sizeof(T) * bb * N);
    if (Rinv) memcpy(Rinv, Ri, sizeof(T) * uu * (N - 1));
    free(Qi); free(Ri); free(aug); free(s_temp); free(AQ); free(BR);
    return 0;
}

int FN(oracle_schur_recover)(const T *Qinv, const T *Rinv, const T *A, const T *B, const T *q, const T *r, const T *lambda,
                             uint32_t nx, uint32_t nu, uint32_t N, T *x, T *u)
{
    if (!Qinv || !Rinv || !A || !B || !q || !r || !lambda || !x || !u || nx == 0 || nu == 0 || N < 2) return -1;
    size_t bb = (size_t)nx * nx, uu = (size_t)nu * nu, xu = (size_t)nx * nu;
    T *tx = (T *)malloc(sizeof(T) * (nx + nu)), *tu = tx + nx;
    if (!tx) return -2;
    for (uint32_t k = 0; k < N; k++) {
        int has_next = k + 1 < N;
        for (uint32_t e = 0; e < nx; e++) {
            T s = (T)0;
            if (has_next) for (uint32_t p = 0; p < nx; p++) s += A[k * bb + (size_t)e * nx + p] * lambda[(size_t)(k + 1) * nx + p];
            tx[e] = (q[(size_t)k * nx + e] + lambda[(size_t)k * nx + e]) - s;
        }
        for (uint32_t e = 0; e < nx; e++) {
            T s = (T)0;
            for (uint32_t p = 0; p < nx; p++) s += Qinv[k * bb + (size_t)p * nx + e] * tx[p];
            x[(size_t)k * nx + e] = -s;
        }
        if (!has_next) continue;
        for (uint32_t e = 0; e < nu; e++) {
            T s = (T)0;
            for (uint32_t p = 0; p < nx; p++) s += B[k * xu + (size_t)e * nx + p] * lambda[(size_t)(k + 1) * nx + p];
            tu[e] = r[(size_t)k * nu + e] - s;
        }
        for (uint32_t e = 0; e < nu; e++) {
            T s = (T)0;
            for (uint32_t p = 0; p < nu; p++) s += Rinv[k * uu + (size_t)p * nu + e] * tu[p];
            u[(size_t)k * nu + e] = -s;
        }
    }
    free(tx);
    return 0;
}

#undef FN
#undef CAT
#undef CAT_
