// fp32 example: nvcc -std=c++17 -I../include pcg_solve.cu -L../gbd_pcg_b200 -lgbdpcg -o pcg.exe
#include "solve_common.h"
int main() { return run_example<float>(2e-4); }
