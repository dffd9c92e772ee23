// pcg_grid_inst.cu -- instantiates the cooperative-grid kernel (pcg_grid.cuh: systems beyond one 16-CTA cluster) for one
// (dtype, nx); compiled once per pair with  -DINST_NX=<nx> -DINST_F64=<0|1>   (see csrc/Makefile)
#include "kernel_table.h"
#include "pcg_grid.cuh"

#if !defined(INST_NX) || !defined(INST_F64)
#error "compile with -DINST_NX=<state size> -DINST_F64=<0|1>"
#endif

namespace gbdpcg {

#if INST_F64
typedef double InstT;
#define INST_DT "f64"
#define INST_RPT 1
#else
typedef float InstT;
#define INST_DT "f32"
#if INST_NX % 2
#define INST_RPT 1
#else
#define INST_RPT 2
#endif
#endif
#define CAT3_(a, b, c) a##b##_##c
#define CAT3(a, b, c) CAT3_(a, b, c)
#if INST_F64
#define INST_TABLE_NAME CAT3(grid_variants_, f64, INST_NX)
#else
#define INST_TABLE_NAME CAT3(grid_variants_, f32, INST_NX)
#endif
#define STR_(x) #x
#define STR(x) STR_(x)
constexpr int RPT = INST_RPT;
constexpr int NX = INST_NX;

// same register budget rule as the cluster kernel (pcg_fused_inst.cu): RPT * 3nx entries per register-resident matrix plus
// working state; warps per sub-partition = 512 / registers
constexpr int regs_needed(int nx, int nmat_reg)
{
    return (int)(sizeof(InstT) / 4) * 3 * nx * RPT * nmat_reg + (sizeof(InstT) == 8 ? (nmat_reg == 2 ? 72 : 96) : 56);
}
constexpr int maxt_for(int nx, int nmat_reg)
{
    int r = (regs_needed(nx, nmat_reg) + 7) / 8 * 8;
    if (r > 255) return 0;
    int t = 128 * (512 / r);
    return t > 1024 ? 1024 : t;
}
constexpr int MAXT_1 = maxt_for(NX, 1) > 0 ? maxt_for(NX, 1) : 32;
constexpr int MAXT_2 = maxt_for(NX, 2);
constexpr int MAXT_2K = MAXT_2 > 0 ? MAXT_2 : 32;

#define VARIANT(PM, MT)                                                                           \
    {NX, INST_F64, RPT, PM, MT, Geo<InstT>::nxp(NX), (const void *)pcg_grid_kernel<InstT, NX, RPT, PM, MT>, \
     "pcg_grid<" INST_DT ",nx=" STR(INST_NX) ",R=" STR(INST_RPT) "," #PM ">",                               \
     (const void *)pcg_grid1_kernel<InstT, NX, RPT, PM, MT>,                                                \
     "pcg_grid1<" INST_DT ",nx=" STR(INST_NX) ",R=" STR(INST_RPT) "," #PM ",one barrier>"}

static const GridVariant g_table[] = {VARIANT(P_NONE, MAXT_1), VARIANT(P_SMEM, MAXT_1), VARIANT(P_REG, MAXT_2K)};

const GridVariant *INST_TABLE_NAME(int *count)
{
    // the P_REG entry is the last one; dropped when two matrices do not fit the register file
    *count = (maxt_for(NX, 1) > 0) ? 3 - ((MAXT_2 > 0) ? 0 : 1) : 0;
    return g_table;
}

}  // namespace gbdpcg
