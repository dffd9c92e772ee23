// pcg_tmem_inst.cu -- instantiations of the tensor-memory batched solver (pcg_tmem.cuh), fp32, even state sizes <= 16.
#include "pcg_tmem.cuh"

namespace gbdpcg {

const void *tmem_kernel_f32(int nx, const char **name)
{
    switch (nx) {
#define CASE(NX) case NX: *name = "pcg_tmem<f32,nx=" #NX ",R=2,2 per SM>"; return (const void *)pcg_tmem_kernel<NX>;
        CASE(8) CASE(10) CASE(12) CASE(14) CASE(16)
#undef CASE
    default: *name = nullptr; return nullptr;
    }
}

}  // namespace gbdpcg
