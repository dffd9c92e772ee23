// constants.cuh -- defaults of the drop-in headers (mirror of reference include/constants.cuh:5-20).
//
// STATE_SIZE / KNOT_POINTS: the reference instantiates its kernel from these macros and silently
// computes garbage when they differ from the run-time arguments (include/interface.cuh:110).  This
// build dispatches on the run-time arguments; the macros are still defined (default 3/3) so that user
// code naming them keeps compiling, but nothing here depends on them.
//
// pcg_constants: same names, types and values.  They are `inline` variables here (the reference
// defines mutable non-inline globals in a header, an ODR violation as soon as two translation units
// include it); they stay mutable because user code may assign to them before building a pcg_config.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>  // dim3

#ifndef STATE_SIZE
#define STATE_SIZE 3
#endif
#ifndef KNOT_POINTS
#define KNOT_POINTS 3
#endif

namespace pcg_constants {
inline uint32_t DEFAULT_MAX_PCG_ITER = 25;
template <typename T>
inline T DEFAULT_EPSILON = static_cast<T>(1e-6);
inline dim3 DEFAULT_GRID(128);   // carried for source compatibility; never used to shape a launch
inline dim3 DEFAULT_BLOCK(64);   // idem (the reference hard-codes its launch shape too, interface.cuh:132)
}  // namespace pcg_constants
