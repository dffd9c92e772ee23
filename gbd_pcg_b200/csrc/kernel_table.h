// kernel_table.h -- registry of compiled register-row kernel variants (one translation unit per (dtype, nx)).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace gbdpcg {

struct RowsVariant {
    int nx;
    int is_f64;
    int clustered;   // 0: one CTA per system (__syncthreads only); 1: thread-block cluster + DSMEM
    int pmode;       // P_NONE / P_REG / P_SMEM
    int maxt;        // __launch_bounds__ of this instantiation = max threads per CTA
    int nxp;         // window stride per block (elements)
    int sb;          // staging stride per block-row (elements)
    const void *fn;  // __global__ entry
    const char *name;
};

// State sizes with a specialised kernel.  Keep in sync with NX_LIST in csrc/Makefile.
#define GBDPCG_FOR_EACH_NX(X) X(2) X(4) X(6) X(8) X(10) X(12) X(14) X(16)
// State sizes that get the fused solver kernel only, in its default single-reduction recurrence (no
// first-generation variants, and the preconditioner builder / Schur kernels use their run-time-nx form): odd sizes
// run one row per thread, sizes above 16 keep Phi^-1 in shared memory (fp64: up to 25, above that a row of S alone
// overflows the register file and the table comes back empty).  Keep in sync with NX_EXTRA in the Makefile.
#define GBDPCG_FOR_EACH_NX_EXTRA(X) \
    X(3) X(5) X(7) X(9) X(11) X(13) X(15) X(17) X(18) X(19) X(20) X(21) X(22) X(23) X(24) X(25) X(26) X(27) X(28) X(29) X(30) X(31) X(32)

#define GBDPCG_DECLARE_TABLE(NX)                               \
    const RowsVariant *rows_variants_f32_##NX(int *count);     \
    const RowsVariant *rows_variants_f64_##NX(int *count);
GBDPCG_FOR_EACH_NX(GBDPCG_DECLARE_TABLE)
#undef GBDPCG_DECLARE_TABLE

// Fused kernel (pcg_fused.cuh): one table per (dtype, nx, recurrence).
struct FusedVariant {
    int nx;
    int is_f64;
    int rpt;         // rows per thread
    int clustered;
    int pmode;
    int fused;       // 1: single-reduction (Chronopoulos-Gear) order, 0: the reference's two-reduction order
    int maxt;
    int nxp;
    const void *fn;
    const char *name;
};
#define GBDPCG_DECLARE_FUSED(NX)                                   \
    const FusedVariant *fused_variants_f32_##NX##_0(int *count);   \
    const FusedVariant *fused_variants_f32_##NX##_1(int *count);   \
    const FusedVariant *fused_variants_f64_##NX##_0(int *count);   \
    const FusedVariant *fused_variants_f64_##NX##_1(int *count);
GBDPCG_FOR_EACH_NX(GBDPCG_DECLARE_FUSED)
#undef GBDPCG_DECLARE_FUSED
#define GBDPCG_DECLARE_FUSED_EXTRA(NX)                             \
    const FusedVariant *fused_variants_f32_##NX##_1(int *count);   \
    const FusedVariant *fused_variants_f64_##NX##_1(int *count);
GBDPCG_FOR_EACH_NX_EXTRA(GBDPCG_DECLARE_FUSED_EXTRA)
#undef GBDPCG_DECLARE_FUSED_EXTRA

// fp64 throughput form of the fused kernel (PMODE = P_TMEM, two rows per thread; pcg_fused_tmem_inst.cu): entries come in pairs
// (one CTA, clustered) per choice of the S columns kept in tensor memory; fused_tmem_sk_f64_<nx>(pair) = that choice.
// Keep in sync with NX_TMEM64 in the Makefile.
#define GBDPCG_FOR_EACH_NX_TMEM64(X) X(8) X(10) X(12) X(14) X(16)
#define GBDPCG_DECLARE_FTMEM(NX)                                   \
    const FusedVariant *fused_tmem_variants_f64_##NX(int *count);  \
    int fused_tmem_sk_f64_##NX(int pair);
GBDPCG_FOR_EACH_NX_TMEM64(GBDPCG_DECLARE_FTMEM)
#undef GBDPCG_DECLARE_FTMEM
// the same form in fp32, four rows per thread (nx 10, 14: two padding rows in the last thread of a block-row).  Keep in sync with NX_TMEM32 in the Makefile.
#define GBDPCG_FOR_EACH_NX_TMEM32(X) X(8) X(10) X(12) X(14) X(16)
#define GBDPCG_DECLARE_FTMEM32(NX)                                 \
    const FusedVariant *fused_tmem_variants_f32_##NX(int *count);  \
    int fused_tmem_sk_f32_##NX(int pair);
GBDPCG_FOR_EACH_NX_TMEM32(GBDPCG_DECLARE_FTMEM32)
#undef GBDPCG_DECLARE_FTMEM32

// Cooperative-grid kernel (pcg_grid.cuh): one table per (dtype, nx).
struct GridVariant {
    int nx;
    int is_f64;
    int rpt;
    int pmode;
    int maxt;
    int nxp;
    const void *fn;
    const char *name;
    const void *fn1;          // pcg_grid1_kernel: one grid barrier per iteration (ghost rows of Phi^-1)
    const char *name1;
};
#define GBDPCG_DECLARE_GRID(NX)                              \
    const GridVariant *grid_variants_f32_##NX(int *count);   \
    const GridVariant *grid_variants_f64_##NX(int *count);
GBDPCG_FOR_EACH_NX(GBDPCG_DECLARE_GRID)
#undef GBDPCG_DECLARE_GRID

}  // namespace gbdpcg
