/*
 * TEST INFRASTRUCTURE ONLY (checker + reported CPU baseline).  Not part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may
 * load liboracle_pcg.so.  See pcg_oracle_impl.h for the reference file:line map.
 *
 * Build:  make -C oracle        (gcc -O3 -march=native -fPIC -shared, no CUDA)
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define T float
#define SUF f32
#define FABS fabsf
#include "pcg_oracle_impl.h"
#undef T
#undef SUF
#undef FABS

#define T double
#define SUF f64
#define FABS fabs
#include "pcg_oracle_impl.h"
#undef T
#undef SUF
#undef FABS

const char *oracle_pcg_version(void) { return "gbd-pcg host oracle 1 (restates reference include/pcg.cuh:118-212)"; }
