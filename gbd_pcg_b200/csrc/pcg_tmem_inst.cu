// pcg_tmem_inst.cu -- instantiations of the tensor-memory batched solvers, fp32, even state sizes <= 16:
// pcg_tmem.cuh (two rows per thread, up to 8 warps per system) and pcg_tmem4.cuh (four rows per thread, up to 4 warps).
#include "pcg_tmem4.cuh"

namespace gbdpcg {

const void *tmem_kernel_f32(int nx, const char **name)
{
    switch (nx) {
#define CASE(NX) case NX: *name = "pcg_tmem<f32,nx=" #NX ",R=2,2 per SM>"; return (const void *)pcg_tmem_kernel<NX>;
        CASE(8) CASE(10) CASE(12) CASE(14) CASE(16)
#undef CASE
    default: *name = nullptr; return nullptr;
    }
}

// Four rows per thread.  `variant` picks where a thread's quads live: (SQT, PQR) = S quads in tensor memory, Phi^-1 quads in
// registers.  Variant 0 is what the planner uses; the others are kept for A/B runs (GBDPCG_TMEM4_VARIANT, scripts/bench_variants.py).
const void *tmem4_kernel_f32(int nx, int variant, const char **name, int *alloc_cols)
{
#define V(NX, ID, SQT, PQR)                                                                 \
    if (nx == NX && variant == ID) {                                                        \
        *name = "pcg_tmem4<f32,nx=" #NX ",R=4,S_tmem=" #SQT ",P_reg=" #PQR ",2 per SM>";    \
        *alloc_cols = T4Geo<NX, SQT, PQR>::alloc_cols;                                      \
        return (const void *)pcg_tmem4_kernel<NX, SQT, PQR>;                                \
    }
    // id 0: what the planner takes for systems of 3..4 warps; id 10: systems of <= 2 warps -- at most 128 tensor-memory columns,
    // so that FOUR CTAs share an SM; the rest: alternatives kept for A/B runs (profiles/tmem4_r2*.txt)
    V(8, 0, 0, 0) V(8, 10, 0, 0)
    V(10, 0, 0, 0) V(10, 10, 0, 0)
    V(12, 0, 3, 3) V(12, 1, 0, 0) V(12, 10, 0, 1)
    V(14, 0, 2, 0) V(14, 1, 4, 0) V(14, 2, 6, 6) V(14, 3, 0, 0) V(14, 10, 0, 3)
    V(16, 0, 6, 2) V(16, 1, 4, 0)
#undef V
    // id 20: the wide form (8 warps, two per lane quarter, one CTA per SM) for systems of 129..256 four-row threads
#define W(NX, SQT, PQR)                                                                     \
    if (nx == NX && variant == 20) {                                                        \
        *name = "pcg_tmem4<f32,nx=" #NX ",R=4,S_tmem=" #SQT ",P_reg=" #PQR ",wide,1 per SM>"; \
        *alloc_cols = 512;                                                                  \
        return (const void *)pcg_tmem4_kernel<NX, SQT, PQR, 2>;                             \
    }
    W(8, 0, 0) W(10, 0, 0) W(12, 3, 3) W(14, 2, 0) W(16, 6, 2)
#undef W
    *name = nullptr;
    *alloc_cols = 0;
    return nullptr;
}

}  // namespace gbdpcg
