// Explicit instantiations: fp16 value with fp32 or fp16 sampling locations / attention weights.
#include "msda_launch.cuh"

namespace msda {
MSDA_DECLARE_TYPED(, __half, float)
MSDA_DECLARE_TYPED(, __half, __half)
}  // namespace msda
