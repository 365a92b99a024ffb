// Stand-in for <ATen/ATen.h> when compiling the reference's ms_deform_im2col_cuda.cuh on its own:
// the kernels in that file use no ATen symbol, only fixed-width integer types.
#pragma once
#include <cstdint>
