// pcg_inst.cu -- instantiates the register-row kernel for one (dtype, nx); compiled once per pair with
//   -DINST_NX=<nx> -DINST_F64=<0|1>     (see csrc/Makefile)
#include "kernel_table.h"
#include "pcg_rows.cuh"

#ifndef INST_NX
#error "compile with -DINST_NX=<state size> -DINST_F64=<0|1>"
#endif

namespace gbdpcg {

#if INST_F64
typedef double InstT;
#define INST_TABLE_NAME_(nx) rows_variants_f64_##nx
#define INST_DT "f64"
#else
typedef float InstT;
#define INST_TABLE_NAME_(nx) rows_variants_f32_##nx
#define INST_DT "f32"
#endif
#define INST_TABLE_NAME(nx) INST_TABLE_NAME_(nx)
#define STR_(x) #x
#define STR(x) STR_(x)

// Register budget -> threads per CTA.  A thread keeps 3*nx matrix entries per register-resident matrix
// plus ~44 (fp32) / ~64 (fp64) 32-bit registers of working state; the SM has 64 K registers.
constexpr int regs_needed(int nx, int nmat_reg)
{
    return (int)(sizeof(InstT) / 4) * 3 * nx * nmat_reg + (sizeof(InstT) == 8 ? 64 : 44);
}
constexpr int maxt_for(int nx, int nmat_reg)
{
    int r = (regs_needed(nx, nmat_reg) + 7) / 8 * 8;
    if (r > 248) return 0;  // does not fit the 255-register ISA limit: variant not built
    int t = 65536 / r / 32 * 32;
    return t > 1024 ? 1024 : t;
}

constexpr int NX = INST_NX;
constexpr int MAXT_1 = maxt_for(NX, 1);                     // S in registers (identity or smem Phi^-1)
constexpr int MAXT_2 = maxt_for(NX, 2);                     // S and Phi^-1 in registers
constexpr int MAXT_2K = MAXT_2 > 0 ? MAXT_2 : 32;           // keeps the template well-formed when unused

#define VARIANT(CL, PM, MT)                                                                                  \
    {NX, INST_F64, CL, PM, MT, Geo<InstT>::nxp(NX), Geo<InstT>::sb(NX),                                       \
     (const void *)pcg_rows_kernel<InstT, NX, (CL != 0), PM, MT>,                                             \
     "pcg_rows<" INST_DT ",nx=" STR(INST_NX) "," #CL "," #PM ">"}

static const RowsVariant g_table[] = {
    VARIANT(0, P_NONE, MAXT_1), VARIANT(1, P_NONE, MAXT_1),
    VARIANT(0, P_SMEM, MAXT_1), VARIANT(1, P_SMEM, MAXT_1),
    VARIANT(0, P_REG, MAXT_2K), VARIANT(1, P_REG, MAXT_2K),
};

const RowsVariant *INST_TABLE_NAME(INST_NX)(int *count)
{
    // the P_REG entries are the last two; drop them when they cannot fit the register file
    *count = (MAXT_2 > 0) ? 6 : 4;
    return g_table;
}

}  // namespace gbdpcg
