// pinv_build.cuh -- builds the preconditioner Phi^-1 from S on the GPU.
//
// The reference host API ships no preconditioner (include/interface.cuh:33-34 prints a warning, :45-46 leaves d_Pinv
// uninitialised) although the tree carries the primitives for one and never calls them: glass::invertMatrix
// (GLASS/src/L3/inv.cuh:6-37, Gauss-Jordan on [A | I] without pivoting), glass::gemm (GLASS/src/L3/gemm.cuh:9-103),
// load_block_bd / store_block_bd (include/utils.cuh:96-161), and the 9*nx^2 scratch term of pcgSharedMemSize
// (include/pcg.cuh:19).  The formulae are those of the paper the reference cites (README.md:66-76):
//     block-Jacobi      Phi^-1_kk      =  S_kk^-1
//     symmetric stair   Phi^-1_{k,k+1} = -S_kk^-1 S_{k,k+1} S_{k+1,k+1}^-1   (right block of block-row k)
//                       Phi^-1_{k,k-1} = -S_kk^-1 S_{k,k-1} S_{k-1,k-1}^-1   (left block of block-row k)
// Output is in the compressed block-tridiagonal layout of the solver (include/pcg.cuh:104-110); padding blocks are zero.
//
// Mapping: a group of nx/CPL lanes owns one nx x nx block; a lane keeps CPL (= 2) columns of it in registers.
//   * inversion: in-place Gauss-Jordan in the reference's arithmetic order (row p scaled by 1/pivot, every other row
//     r updated as a[r] -= (col_p[r] * pvInv) * a[p], inv.cuh:27-33); the pivot column is broadcast through a
//     16-float shared scratch line per group, __syncwarp only.
//   * products: out[:, j] = -(D_k^-1 (S_off (D_kn^-1[:, j]))): two register mat-vecs per lane whose matrix operand is
//     streamed from shared memory with 16-byte loads that every lane of the group shares (broadcast).
// A CTA stages a chunk of block-rows of S with coalesced 16-byte loads, rewrites the image in place into Phi^-1 and
// writes it back with coalesced 16-byte stores: every byte of S is read once and every byte of Phi^-1 written once
// (plus two halo diagonal blocks per chunk when a system is cut into several chunks).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gbdpcg {

#ifndef PINV_L2_PREFETCH
#define PINV_L2_PREFETCH 1
#endif

enum { PINV_IDENTITY = 0, PINV_JACOBI = 1, PINV_STAIR = 2 };

struct PinvArgs {
    const void *S;
    void *P;
    unsigned long long mat_stride;   // elements between consecutive problems
    uint32_t N, batch, chunk, nchunks;
    int kind;
    int use_tma;                     // host decided: S, P and the problem stride are 16-byte aligned
};

// ---- TMA engine (1-D bulk copies) + mbarrier, used for the chunk image in both directions ----
__device__ __forceinline__ uint32_t pv_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pv_mbar_init(uint32_t mbar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void pv_mbar_expect_tx(uint32_t mbar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pv_mbar_wait(uint32_t mbar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PV_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra PV_DONE_%=;\n"
        "bra PV_WAIT_%=;\n"
        "PV_DONE_%=:\n"
        "}\n" ::"r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void pv_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void pv_bulk_s2g(void *dst, uint32_t src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void pv_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void pv_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <typename T> struct PVec;
template <> struct PVec<float> { typedef float4 type; static constexpr int V = 4; };
template <> struct PVec<double> { typedef double2 type; static constexpr int V = 2; };

template <typename T, int V> __device__ __forceinline__ void unpack_vec(const typename PVec<T>::type &q, T (&x)[4])
{
    if constexpr (V == 4) { x[0] = q.x; x[1] = q.y; x[2] = q.z; x[3] = q.w; }
    else { x[0] = q.x; x[1] = q.y; x[2] = (T)0; x[3] = (T)0; }
}

// Y = M V for a column-major NX x NX block M in shared memory and C columns V held by the lane (flat 16-byte loads
// shared by the whole group; every element of M that reaches a register is used for C multiply-adds)
template <typename T, int NX, int C>
__device__ __forceinline__ void block_mv_bcast(const T *__restrict__ M, const T (&v)[C][NX], T (&y)[C][NX])
{
    constexpr int V = PVec<T>::V;
#pragma unroll
    for (int c = 0; c < C; c++)
#pragma unroll
        for (int r = 0; r < NX; r++) y[c][r] = (T)0;
    if constexpr ((NX * NX) % V == 0) {
#pragma unroll
        for (int e = 0; e < NX * NX; e += V) {
            const typename PVec<T>::type q = *reinterpret_cast<const typename PVec<T>::type *>(M + e);
            T x[4];
            unpack_vec<T, V>(q, x);
#pragma unroll
            for (int u = 0; u < V; u++)
#pragma unroll
                for (int c = 0; c < C; c++) y[c][(e + u) % NX] += x[u] * v[c][(e + u) / NX];
        }
    } else {
#pragma unroll
        for (int e = 0; e < NX * NX; e++)
#pragma unroll
            for (int c = 0; c < C; c++) y[c][e % NX] += M[e] * v[c][e / NX];
    }
}

// In-place Gauss-Jordan inversion of one NX x NX block spread over the NX/CPL lanes of a group: lane `li` holds columns
// li*CPL .. li*CPL+CPL-1 in `col`.  Per pivot p the owner of column p publishes the scaled pivot column
// m[r] = col_p[r] * pvInv (m[p] = pvInv) through the group's 16-element shared line `scr`; every lane then applies
// a[r] -= m[r] * a[p], a[p] *= pvInv to its columns.  The owner first turns its column p into e_p, so the same update leaves
// the image of the augmented identity column there: exactly the values of Gauss-Jordan on [A | I] in the arithmetic order
// of glass::invertMatrix (reference GLASS/src/L3/inv.cuh:27-33).  All lanes of the warp must call it (__syncwarp inside);
// lanes with valid == false compute on dummies.
// `nsteps` (warp-uniform, <= NX): pivots to run; a block padded with an identity tail (schur_form.cuh pads R_k) needs
// only its leading nsteps pivots -- the remaining ones would multiply by exactly 1 and subtract exactly 0.
template <typename T, int NX, int CPL>
__device__ __forceinline__ void gj_invert_cols(T (&col)[CPL][NX], const bool valid, const int li, T *scr, const int nsteps = NX)
{
    constexpr int V = PVec<T>::V;
#pragma unroll
    for (int p = 0; p < NX; p++) {
        if (p >= nsteps) break;
        if (valid && li == p / CPL) {
            const T pv_inv = (T)1 / col[p % CPL][p];
#pragma unroll
            for (int r = 0; r < NX; r++) {
                scr[r] = (r == p) ? pv_inv : col[p % CPL][r] * pv_inv;
                col[p % CPL][r] = (r == p) ? (T)1 : (T)0;
            }
        }
        __syncwarp();
        T m[NX + 4];
#pragma unroll
        for (int r = 0; r < NX; r += V) {
            T x[4];
            unpack_vec<T, V>(*reinterpret_cast<const typename PVec<T>::type *>(scr + r), x);
#pragma unroll
            for (int u = 0; u < V; u++) m[r + u] = x[u];
        }
#pragma unroll
        for (int c = 0; c < CPL; c++) {
            const T prv = col[c][p];
#pragma unroll
            for (int r = 0; r < NX; r++) col[c][r] = (r == p) ? prv * m[p] : col[c][r] - m[r] * prv;
        }
        __syncwarp();
    }
}

template <typename T>
__device__ __forceinline__ void copy_coalesced(T *__restrict__ dst, const T *__restrict__ src, int count, bool src_global)
{
    const int t = threadIdx.x, nthr = blockDim.x;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0 && (count * (int)sizeof(T)) % 16 == 0) {
        const int total = count * (int)sizeof(T) / 16;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4 *d4 = reinterpret_cast<uint4 *>(dst);
        constexpr int U = 4;
        for (int base = 0; base < total; base += U * nthr) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < total) v[u] = src_global ? __ldg(s4 + q) : s4[q]; }
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < total) d4[q] = v[u]; }
        }
    } else {
        constexpr int U = 4;                                   // loads first, then stores: U requests in flight per thread
        for (int base = 0; base < count; base += U * nthr) {
            T v[U];
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < count) v[u] = src[q]; }
#pragma unroll
            for (int u = 0; u < U; u++) { const int q = base + u * nthr + t; if (q < count) dst[q] = v[u]; }
        }
    }
}

// Asynchronous global -> shared copy of `count` elements with cp.async (LDGSTS): 16-byte pieces when both sides are
// 16-byte aligned, element-sized pieces otherwise.  Nothing is waited for here: issue several, then async_copy_wait().
template <typename T>
__device__ __forceinline__ void async_copy(T *__restrict__ dst, const T *__restrict__ src, int count)
{
    const int t = threadIdx.x, nthr = blockDim.x;
    if (count <= 0) return;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && (pv_smem_u32(dst) & 15u) == 0 && (count * (int)sizeof(T)) % 16 == 0) {
        const int total = count * (int)sizeof(T) / 16;
        const uint32_t d0 = pv_smem_u32(dst);
        const char *s0 = reinterpret_cast<const char *>(src);
        for (int q = t; q < total; q += nthr)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * (uint32_t)q), "l"(s0 + 16 * (size_t)q) : "memory");
    } else {
        const uint32_t d0 = pv_smem_u32(dst);
        for (int q = t; q < count; q += nthr) {
            if constexpr (sizeof(T) == 4)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4u * (uint32_t)q), "l"(src + q) : "memory");
            else
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + 8u * (uint32_t)q), "l"(src + q) : "memory");
        }
    }
}
__device__ __forceinline__ void async_copy_wait()
{
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// columns of a block one lane owns during the inversion (2 halves the per-pivot overhead per block; 1 for tiny blocks)
__host__ __device__ constexpr int pinv_cpl(int nx) { return (nx >= 4 && nx % 2 == 0) ? 2 : 1; }

// shared memory: [image: chunk * 3 NX^2] [dinv: (chunk + 2) * NX^2] [pivot-column scratch: groups * 16]
template <typename T>
__host__ __device__ inline size_t pinv_smem_bytes(int nx, int chunk, int threads)
{
    const int groups = (threads / 32) * (32 / (nx / pinv_cpl(nx)));
    size_t e = (size_t)chunk * 3 * nx * nx + (size_t)(chunk + 2) * nx * nx + (size_t)groups * 16;
    return e * sizeof(T) + 48;       // + mbarrier slot in front
}

template <typename T, int NX>
__global__ void __launch_bounds__(256) pinv_build_kernel(const PinvArgs a)
{
    constexpr int BB = NX * NX, BSZ = 3 * BB;
    extern __shared__ __align__(16) unsigned char pinv_smem_raw[];
    const uint32_t mbar = pv_smem_u32(pinv_smem_raw);
    T *img = reinterpret_cast<T *>(pinv_smem_raw + 16);
    const int chunk = (int)a.chunk, N = (int)a.N;
    T *dinv = img + (size_t)chunk * BSZ;
    T *scr_all = dinv + (size_t)(chunk + 2) * BB;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    // lane mapping: LPB lanes per block, CPL columns of it per lane
    constexpr int CPL = pinv_cpl(NX), LPB = NX / CPL, GPW1 = 32 / LPB;
    const int gl1 = lane / LPB, li = lane - gl1 * LPB;
    const bool lane_on1 = gl1 < GPW1;
    const int group1 = warp * GPW1 + gl1;
    const int ngroups1 = (blockDim.x >> 5) * GPW1;
    T *scr1 = scr_all + (lane_on1 ? group1 : 0) * 16;

    const T *Sg = static_cast<const T *>(a.S);
    T *Pg = static_cast<T *>(a.P);
    const uint32_t items = a.batch * a.nchunks;
    const bool tma = a.use_tma != 0;
    uint32_t phase = 0;
    if (tma && t == 0) pv_mbar_init(mbar, 1);
    __syncthreads();

    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const uint32_t prob = item / a.nchunks;
        const int k0 = (int)(item - prob * a.nchunks) * chunk;
        const int cnt = min(chunk, N - k0);
        const T *Sp = Sg + (size_t)prob * a.mat_stride;
        T *Pp = Pg + (size_t)prob * a.mat_stride;

        if (a.kind == PINV_IDENTITY) {
            for (int e = t; e < cnt * BSZ; e += blockDim.x) {
                const int w = e % BSZ, c = w / NX, r = w - c * NX;
                img[e] = (c == NX + r) ? (T)1 : (T)0;
            }
        } else {
            if (tma) {
                // one thread: the previous item's bulk store must have finished READING the image, then the TMA engine
                // fetches this item's chunk of S; everybody waits on the mbarrier
                if (t == 0) {
                    pv_bulk_wait_read();
                    const uint32_t bytes = (uint32_t)(cnt * BSZ * (int)sizeof(T));
                    pv_mbar_expect_tx(mbar, bytes);
                    pv_bulk_g2s(pv_smem_u32(img), Sp + (size_t)k0 * BSZ, bytes, mbar);
                }
                pv_mbar_wait(mbar, phase);
                phase ^= 1;
#if PINV_L2_PREFETCH
                // the image is rewritten in place, so the next item's copy cannot start before this item's store has read it: pull the
                // next chunk of S into L2 meanwhile (its copy then waits on L2, not on HBM)
                if (t == 32 && item + gridDim.x < items) {
                    const uint32_t nitem = item + gridDim.x, nprob = nitem / a.nchunks;
                    const int nk0 = (int)(nitem - nprob * a.nchunks) * chunk, ncnt = min(chunk, N - nk0);
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(Sg + (size_t)nprob * a.mat_stride + (size_t)nk0 * BSZ),
                                 "r"((uint32_t)(ncnt * BSZ * (int)sizeof(T))) : "memory");
                }
#endif
            } else {
                copy_coalesced<T>(img, Sp + (size_t)k0 * BSZ, cnt * BSZ, true);
                __syncthreads();
            }

            // ---- phase 1: D_k^-1 for k = k0-1 .. k0+cnt (halo blocks come straight from global memory) ----
            // CPL columns per lane, LPB lanes per block.  Per pivot p the owner of column p publishes the scaled pivot
            // column m[r] = col_p[r] * pvInv (m[p] = pvInv); every lane then applies a[r] -= m[r] * a[p], a[p] *= pvInv
            // to its columns.  The owner first turns its column p into e_p, so the same update leaves the image of
            // the augmented identity column there: exactly the values of Gauss-Jordan on [A | I] (inv.cuh:27-33).
            // task list = the knots whose inverse is needed and exists (halo slots 0 and cnt+1 only for the stair kind)
            const int i_first = (a.kind == PINV_STAIR && k0 > 0) ? 0 : 1;
            const int i_last = (a.kind == PINV_STAIR && k0 + cnt < N) ? cnt + 1 : cnt;
            const int ninv = i_last - i_first + 1;
            for (int base = 0; base < ninv; base += ngroups1) {
                const int i = i_first + base + group1;
                const int k = k0 - 1 + i;
                const bool inside = i >= 1 && i <= cnt;
                const bool valid = lane_on1 && i <= i_last;
                if (!__any_sync(0xffffffffu, valid)) continue;          // warp-uniform: this warp has no block in this round
                T col[CPL][NX];
#pragma unroll
                for (int c = 0; c < CPL; c++)
#pragma unroll
                    for (int r = 0; r < NX; r++) col[c][r] = (r == li * CPL + c) ? (T)1 : (T)0;
                if (valid) {
                    const T *src = (inside ? img + (size_t)(i - 1) * BSZ + BB : Sp + (size_t)k * BSZ + BB) + li * CPL * NX;
#pragma unroll
                    for (int c = 0; c < CPL; c++)
#pragma unroll
                        for (int r = 0; r < NX; r++) col[c][r] = src[c * NX + r];
                }
                gj_invert_cols<T, NX, CPL>(col, valid, li, scr1);
                if (valid) {
                    T *d = dinv + (size_t)i * BB + li * CPL * NX;
#pragma unroll
                    for (int c = 0; c < CPL; c++)
#pragma unroll
                        for (int r = 0; r < NX; r++) d[c * NX + r] = col[c][r];
                    if (inside) {
                        T *o = img + (size_t)(i - 1) * BSZ + BB + li * CPL * NX;
#pragma unroll
                        for (int c = 0; c < CPL; c++)
#pragma unroll
                            for (int r = 0; r < NX; r++) o[c * NX + r] = col[c][r];
                    }
                }
            }
            __syncthreads();

            // ---- phase 2: off-diagonal blocks, rewritten in place (same lane mapping: CPL columns per lane) ----
            const int ntask = 2 * cnt;
            for (int base = 0; base < ntask; base += ngroups1) {
                const int i = base + group1;
                const bool valid = lane_on1 && i < ntask;
                const int lb = valid ? (i >> 1) : 0, side = i & 1;       // side 0: left block (k, k-1); 1: right block (k, k+1)
                const int k = k0 + lb, kn = side ? k + 1 : k - 1;
                const bool real = valid && a.kind == PINV_STAIR && kn >= 0 && kn < N;
                T *blk = img + (size_t)lb * BSZ + (side ? 2 * BB : 0);
                T out[CPL][NX];
#pragma unroll
                for (int c = 0; c < CPL; c++)
#pragma unroll
                    for (int r = 0; r < NX; r++) out[c][r] = (T)0;
                if (real) {
                    T v[CPL][NX], w[CPL][NX];
                    const T *dn = dinv + (size_t)(kn - k0 + 1) * BB + li * CPL * NX;
#pragma unroll
                    for (int c = 0; c < CPL; c++)
#pragma unroll
                        for (int r = 0; r < NX; r++) v[c][r] = dn[c * NX + r];
                    block_mv_bcast<T, NX, CPL>(blk, v, w);                              // w = S_off (D_kn^-1)[:, cols]
                    block_mv_bcast<T, NX, CPL>(dinv + (size_t)(lb + 1) * BB, w, out);   // out = D_k^-1 w
#pragma unroll
                    for (int c = 0; c < CPL; c++)
#pragma unroll
                        for (int r = 0; r < NX; r++) out[c][r] = -out[c][r];
                }
                __syncwarp();      // every lane of the group has finished reading S_off before it is overwritten
                if (valid) {
#pragma unroll
                    for (int c = 0; c < CPL; c++)
#pragma unroll
                        for (int r = 0; r < NX; r++) blk[(li * CPL + c) * NX + r] = out[c][r];
                }
            }
        }
        if (tma) {
            pv_fence_async();          // generic-proxy writes of the image become visible to the TMA engine
            __syncthreads();
            if (t == 0) pv_bulk_s2g(Pp + (size_t)k0 * BSZ, pv_smem_u32(img), (uint32_t)(cnt * BSZ * (int)sizeof(T)));
            // the image is not touched again before thread 0 has waited for the store to finish reading it (loop top)
            if (a.kind == PINV_IDENTITY) { if (t == 0) pv_bulk_wait_read(); __syncthreads(); }
        } else {
            __syncthreads();
            copy_coalesced<T>(Pp + (size_t)k0 * BSZ, img, cnt * BSZ, false);
            __syncthreads();
        }
    }
    if (tma && t == 0) pv_bulk_wait_read();   // shared memory must outlive the last bulk store
}

// Runtime-nx fallback (odd or large state sizes): one CTA per (problem, block-row); cooperative in-place Gauss-Jordan of
// the up to three diagonal blocks involved, same arithmetic order as above.
// shared memory: inv[3][nx*nx] | tmp[nx*nx] | colv[nx] | rowv[nx]
template <typename T>
__host__ __device__ inline size_t pinv_generic_smem_bytes(uint32_t nx) { return ((size_t)4 * nx * nx + 2 * nx) * sizeof(T) + 16; }

template <typename T>
__global__ void __launch_bounds__(128) pinv_build_generic_kernel(const PinvArgs a, const uint32_t nx)
{
    extern __shared__ __align__(16) unsigned char pinv_smem_raw[];
    T *inv = reinterpret_cast<T *>(pinv_smem_raw);
    const int bb = (int)(nx * nx), bsz = 3 * bb, N = (int)a.N;
    T *tmp = inv + 3 * bb, *colv = tmp + bb, *rowv = colv + nx;
    const int t = threadIdx.x, nthr = blockDim.x;
    const uint32_t items = a.batch * a.N;
    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
        const uint32_t prob = item / a.N;
        const int k = (int)(item - prob * a.N);
        const T *Sp = static_cast<const T *>(a.S) + (size_t)prob * a.mat_stride;
        T *Pr = static_cast<T *>(a.P) + (size_t)prob * a.mat_stride + (size_t)k * bsz;
        if (a.kind == PINV_IDENTITY) {
            for (int e = t; e < bsz; e += nthr) { const int c = e / (int)nx, r = e - c * (int)nx; Pr[e] = (c == (int)nx + r) ? (T)1 : (T)0; }
            continue;
        }
        const int nblk = (a.kind == PINV_STAIR) ? 3 : 1;           // which diagonal blocks are needed: k (slot 1), k-1 (0), k+1 (2)
        for (int s = 0; s < 3; s++) {
            const int kk = k - 1 + s;
            if ((nblk == 1 && s != 1) || kk < 0 || kk >= N) continue;
            T *A = inv + s * bb;
            for (int e = t; e < bb; e += nthr) A[e] = Sp[(size_t)kk * bsz + bb + e];
            __syncthreads();
            for (int p = 0; p < (int)nx; p++) {
                for (int e = t; e < (int)nx; e += nthr) { colv[e] = A[p * (int)nx + e]; rowv[e] = A[e * (int)nx + p]; }
                __syncthreads();
                const T pv_inv = (T)1 / colv[p];
                for (int e = t; e < bb; e += nthr) {
                    const int c = e / (int)nx, r = e - c * (int)nx;
                    const T m = colv[r] * pv_inv;
                    T v;
                    if (r == p) v = (c == p) ? pv_inv : rowv[c] * pv_inv;
                    else v = (c == p) ? -m : A[e] - m * rowv[c];
                    A[e] = v;
                }
                __syncthreads();
            }
        }
        for (int e = t; e < bb; e += nthr) Pr[bb + e] = inv[bb + e];
        for (int side = 0; side < 2; side++) {
            const int kn = side ? k + 1 : k - 1;
            T *dst = Pr + (side ? 2 * bb : 0);
            if (a.kind != PINV_STAIR || kn < 0 || kn >= N) {
                for (int e = t; e < bb; e += nthr) dst[e] = (T)0;
                continue;
            }
            const T *So = Sp + (size_t)k * bsz + (side ? 2 * bb : 0);
            const T *Dn = inv + (side ? 2 : 0) * bb, *Dk = inv + bb;
            for (int e = t; e < bb; e += nthr) {                  // tmp = S_off * D_kn^-1
                const int c = e / (int)nx, r = e - c * (int)nx;
                T s = (T)0;
                for (int q = 0; q < (int)nx; q++) s += So[q * (int)nx + r] * Dn[c * (int)nx + q];
                tmp[e] = s;
            }
            __syncthreads();
            for (int e = t; e < bb; e += nthr) {                  // dst = -(D_k^-1 * tmp)
                const int c = e / (int)nx, r = e - c * (int)nx;
                T s = (T)0;
                for (int q = 0; q < (int)nx; q++) s += Dk[q * (int)nx + r] * tmp[c * (int)nx + q];
                dst[e] = -s;
            }
            __syncthreads();
        }
        __syncthreads();
    }
}

}  // namespace gbdpcg
