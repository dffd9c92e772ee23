// fp64 example: nvcc -std=c++17 -I../include pcg_solve_dp.cu -L../gbd_pcg_b200 -lgbdpcg -o pcg_dp.exe
#include "solve_common.h"
int main() { return run_example<double>(1e-9); }
