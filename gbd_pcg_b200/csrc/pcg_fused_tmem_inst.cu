// pcg_fused_tmem_inst.cu -- instantiates the tensor-memory form of the fused kernel (pcg_fused.cuh, PMODE = P_TMEM): Phi^-1 rows and the
// last SK panel columns of the S rows in tensor memory; fp64 with two rows per thread, fp32 (-DINST_F32=1, nx a multiple of 4) with four --
// either way 12 nx words of S and 4 (3 nx + SK) tensor-memory columns per thread.  Compiled once per (dtype, state size) with
// -DINST_NX=<nx> [-DINST_F32=1] (see csrc/Makefile).
#include "kernel_table.h"
#include "pcg_fused.cuh"

#if !defined(INST_NX)
#error "compile with -DINST_NX=<state size>"
#endif

namespace gbdpcg {

#ifndef INST_F32
#define INST_F32 0
#endif
#if INST_F32
typedef float InstT;
constexpr int RPT = 4;
#define INST_DT "f32"
#define INST_RS "4"
#define TABLE_NAME CAT2(fused_tmem_variants_f32_, INST_NX)
#define SK_NAME CAT2(fused_tmem_sk_f32_, INST_NX)
#else
typedef double InstT;
constexpr int RPT = 2;
#define INST_DT "f64"
#define INST_RS "2"
#define TABLE_NAME CAT2(fused_tmem_variants_f64_, INST_NX)
#define SK_NAME CAT2(fused_tmem_sk_f64_, INST_NX)
#endif
#define CAT2_(a, b) a##b
#define CAT2(a, b) CAT2_(a, b)
#define STR_(x) #x
#define STR(x) STR_(x)
constexpr int NX = INST_NX;
constexpr int MAXT_BIG = 256;  // 8 warps: two per tensor-memory lane quarter
// Panel columns of S that go to tensor memory beside Phi^-1: what the register file cannot hold (2 rows x 3nx doubles of S are 12nx
// registers; vectors, accumulators, window pairs and two tensor-memory chunks need about 130 more: nx 14 with 12 columns -> 250
// registers, no spills; with 8 -> 255 and 48 bytes of spills; one chunk more measured equal or up to 4 % faster on every state
// size, profiles/tmem64_r2r.txt).  A/B alternative in the second pair.
#ifndef INST_SK_A
#define INST_SK_A (((12 * INST_NX > 120) ? ((12 * INST_NX - 120 + 15) / 16 * 4) : 0) + 4)
#endif
#ifndef INST_SK_B
#define INST_SK_B ((INST_SK_A) + 4)
#endif
constexpr int SK_A = INST_SK_A, SK_B = INST_SK_B;
// CTAs of at most 128 threads (one warp per lane quarter) allocate 256 columns, so that TWO of them share an SM: Phi^-1 (12 nx columns)
// plus the S tail must fit those 256
constexpr int sk_small(int sk) { return (12 * NX + 4 * sk <= 256) ? sk : (256 - 12 * NX) / 16 * 4; }
constexpr int SKS_A = sk_small(SK_A), SKS_B = sk_small(SK_B);
constexpr int MAXT_SMALL = 128;
// (nx 16: two warps of a lane quarter would need 2 x 288 columns -- its "big" kernels are 128-thread ones too, alone on their SM)
constexpr int MAXT = (2 * (12 * NX + 4 * SK_B) <= 512) ? MAXT_BIG : MAXT_SMALL;

#define VARIANT(CL, MT, SKV, SKNAME)                                                                          \
    {NX, !INST_F32, RPT, CL, P_TMEM, 1, MT, Geo<InstT>::nxp(NX),                                                \
     (const void *)pcg_fused_kernel<InstT, NX, RPT, P_TMEM, true, (CL != 0), MT, SKV>,                        \
     "pcg_fused<" INST_DT ",nx=" STR(INST_NX) ",R=" INST_RS "," #CL ",P_TMEM,S_tmem=" SKNAME ">"}

// [0..3]: up to 256 threads, one CTA per SM; [4..7]: up to 128 threads, two CTAs per SM.  Each: (one CTA, clustered) x (choice a, b).
static const FusedVariant g_table[] = {
    VARIANT(0, MAXT, SK_A, "a"), VARIANT(1, MAXT, SK_A, "a"),
    VARIANT(0, MAXT, SK_B, "b"), VARIANT(1, MAXT, SK_B, "b"),
    VARIANT(0, MAXT_SMALL, SKS_A, "a,2 per SM"), VARIANT(1, MAXT_SMALL, SKS_A, "a,2 per SM"),
    VARIANT(0, MAXT_SMALL, SKS_B, "b,2 per SM"), VARIANT(1, MAXT_SMALL, SKS_B, "b,2 per SM"),
};

const FusedVariant *TABLE_NAME(int *count)
{
    *count = (int)(sizeof g_table / sizeof g_table[0]);
    return g_table;
}
// columns of S in tensor memory: entry pair 0 / 1 of the 256-thread kernels, 2 / 3 of the 128-thread ones
int SK_NAME(int pair) { return pair == 0 ? SK_A : pair == 1 ? SK_B : pair == 2 ? SKS_A : SKS_B; }

}  // namespace gbdpcg
