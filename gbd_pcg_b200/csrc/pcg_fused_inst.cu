// pcg_fused_inst.cu -- instantiates the fused kernel (pcg_fused.cuh) for one (dtype, nx, recurrence); compiled
// once per triple with  -DINST_NX=<nx> -DINST_F64=<0|1> -DINST_FUSED=<0|1>   (see csrc/Makefile)
#include "kernel_table.h"
#include "pcg_fused.cuh"

#if !defined(INST_NX) || !defined(INST_F64) || !defined(INST_FUSED)
#error "compile with -DINST_NX=<state size> -DINST_F64=<0|1> -DINST_FUSED=<0|1>"
#endif

namespace gbdpcg {

#if INST_F64
typedef double InstT;
#define INST_DT "f64"
#define INST_RPT 1       // rows per thread: a double row of S and Phi^-1 already fills the register file
#else
typedef float InstT;
#define INST_DT "f32"
#if INST_NX % 2
#define INST_RPT 1       // odd state size: one row per thread
#else
#define INST_RPT 2       // two rows per thread: window loads, shuffles and scalar work are shared by both rows
#endif
#endif
#define CAT4_(a, b, c, d) a##b##_##c##_##d
#define CAT4(a, b, c, d) CAT4_(a, b, c, d)
#if INST_F64
#define INST_TABLE_NAME CAT4(fused_variants_, f64, INST_NX, INST_FUSED)
#else
#define INST_TABLE_NAME CAT4(fused_variants_, f32, INST_NX, INST_FUSED)
#endif
constexpr int RPT = INST_RPT;
#define STR_(x) #x
#define STR(x) STR_(x)

// Register budget -> threads per CTA.  A thread keeps RPT * 3*nx matrix entries per register-resident matrix plus
// working state; the SM has 64 K registers, a thread at most 255.
constexpr int regs_needed(int nx, int nmat_reg)
{
    return (int)(sizeof(InstT) / 4) * 3 * nx * RPT * nmat_reg + (sizeof(InstT) == 8 ? (nmat_reg == 2 ? 72 : 96) : 56);
}
// The register file is split over the 4 SM sub-partitions (16 K registers each) and warps are dealt to them
// round-robin, so what bounds registers per thread is warps per sub-partition: 512 / regs, whole warps.
constexpr int maxt_for(int nx, int nmat_reg)
{
    int r = (regs_needed(nx, nmat_reg) + 7) / 8 * 8;
    if (r > 255) return 0;
    int wps = 512 / r;   // warps per sub-partition that still leave r registers per thread
    int t = 128 * wps;
    return t > 1024 ? 1024 : t;
}

constexpr int NX = INST_NX;
constexpr int MAXT_1 = maxt_for(NX, 1) > 0 ? maxt_for(NX, 1) : 32;
constexpr int MAXT_2 = maxt_for(NX, 2);
constexpr int MAXT_2K = MAXT_2 > 0 ? MAXT_2 : 32;
constexpr bool FUSED = INST_FUSED != 0;

#define VARIANT(CL, PM, MT)                                                                                   \
    {NX, INST_F64, RPT, CL, PM, INST_FUSED, MT, Geo<InstT>::nxp(NX),                                            \
     (const void *)pcg_fused_kernel<InstT, NX, RPT, PM, FUSED, (CL != 0), MT>,                                 \
     "pcg_fused<" INST_DT ",nx=" STR(INST_NX) ",R=" STR(INST_RPT) "," #CL "," #PM "," STR(INST_FUSED) ">"}

static const FusedVariant g_table[] = {
    VARIANT(0, P_NONE, MAXT_1), VARIANT(1, P_NONE, MAXT_1),
    VARIANT(0, P_SMEM, MAXT_1), VARIANT(1, P_SMEM, MAXT_1),
    VARIANT(0, P_REG, MAXT_2K), VARIANT(1, P_REG, MAXT_2K),
};

const FusedVariant *INST_TABLE_NAME(int *count)
{
    // the row-layout P_REG entries are the last two; dropped when they cannot fit the register file.  A state size whose
    // S rows alone overflow it has no usable entry at all (the planner then takes the run-time-nx kernel).
    *count = (maxt_for(NX, 1) > 0) ? (int)(sizeof g_table / sizeof g_table[0]) - ((MAXT_2 > 0) ? 0 : 2) : 0;
    return g_table;
}

}  // namespace gbdpcg
