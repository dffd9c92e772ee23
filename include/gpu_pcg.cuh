// gpu_pcg.cuh -- umbrella header users include (same name as the reference's include/gpu_pcg.cuh).
// Link with -lgbdpcg (gbd_pcg_b200/libgbdpcg.so).  See INTEGRATION.md.
#pragma once
#include "interface.cuh"
