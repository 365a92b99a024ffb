// ref_wrap.cu -- TEST INFRASTRUCTURE.  Thin extern "C" front-end over the REFERENCE's own CUDA
// kernels, compiled from the reference sources where they lie (nothing is copied into this repo):
//   #include "cuda/ms_deform_im2col_cuda.cuh"  resolves to
//   /root/reference/src/trackformer/models/ops/src/cuda/ms_deform_im2col_cuda.cuh
// via -I on the nvcc command line (oracle/ref_cuda/build_ref.py).  The three launchers called here
// are that file's ms_deformable_im2col_cuda (:380-410), ms_deformable_col2im_cuda (:412-452) and
// ms_deformable_col2im_coord_cuda (:454-496).  Output: oracle/_ref/libmsda_ref_cuda.so (git-ignored).
// Used as (a) a second oracle at full BASELINE sizes on the GPU box and (b) the same-box GPU baseline
// ("the kernel to beat").  The product never links or loads it.
#include "cuda/ms_deform_im2col_cuda.cuh"

#define DEFINE_REF(T, SFX)                                                                              \
    extern "C" int msda_ref_im2col_##SFX(void* stream, const T* value, const int64_t* shapes,           \
                                         const int64_t* level_start, const T* loc, const T* attn,       \
                                         int N, int S, int M, int D, int L, int Lq, int P, T* columns)  \
    {                                                                                                   \
        ms_deformable_im2col_cuda<T>((cudaStream_t)stream, value, shapes, level_start, loc, attn, N, S, \
                                     M, D, L, Lq, P, columns);                                          \
        return (int)cudaGetLastError();                                                                 \
    }                                                                                                   \
    extern "C" int msda_ref_col2im_##SFX(void* stream, const T* grad_out, const int64_t* shapes,        \
                                         const int64_t* level_start, const T* loc, const T* attn,       \
                                         int N, int S, int M, int D, int L, int Lq, int P, T* grad_value) \
    {                                                                                                   \
        ms_deformable_col2im_cuda<T>((cudaStream_t)stream, grad_out, shapes, level_start, loc, attn, N, \
                                     S, M, D, L, Lq, P, grad_value);                                    \
        return (int)cudaGetLastError();                                                                 \
    }                                                                                                   \
    extern "C" int msda_ref_col2im_coord_##SFX(void* stream, const T* grad_out, const T* value,         \
                                               const int64_t* shapes, const int64_t* level_start,       \
                                               const T* loc, const T* attn, int N, int S, int M, int D, \
                                               int L, int Lq, int P, T* grad_loc, T* grad_attn)         \
    {                                                                                                   \
        ms_deformable_col2im_coord_cuda<T>((cudaStream_t)stream, grad_out, value, shapes, level_start,  \
                                           loc, attn, N, S, M, D, L, Lq, P, grad_loc, grad_attn);       \
        return (int)cudaGetLastError();                                                                 \
    }

DEFINE_REF(float, f32)
DEFINE_REF(double, f64)
