// Explicit instantiations: bf16 value with fp32 or bf16 sampling locations / attention weights.
#include "msda_launch.cuh"

namespace msda {
MSDA_DECLARE_TYPED(, __nv_bfloat16, float)
MSDA_DECLARE_TYPED(, __nv_bfloat16, __nv_bfloat16)
}  // namespace msda

#ifdef MSDA_DENSE_TIMING
// development builds only: read and clear the dense-window kernel's per-phase cycle counters (this translation
// unit's copy of the device symbol: the bf16 kernels live here)
extern "C" int msda_b200_debug_dense_phase_cycles(unsigned long long* out16)
{
    unsigned long long zero[16] = {0};
    cudaError_t e = cudaMemcpyFromSymbol(out16, msda::dense::dense_phase_cycles, sizeof(zero));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(msda::dense::dense_phase_cycles, zero, sizeof(zero));
    return (int)e;
}
#endif
