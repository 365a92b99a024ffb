// Explicit instantiations: bf16 value with fp32 or bf16 sampling locations / attention weights.
#include "msda_launch.cuh"

namespace msda {
MSDA_DECLARE_TYPED(, __nv_bfloat16, float)
MSDA_DECLARE_TYPED(, __nv_bfloat16, __nv_bfloat16)
}  // namespace msda
