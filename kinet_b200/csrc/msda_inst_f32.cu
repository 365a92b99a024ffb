// Explicit instantiations: fp32 value (vector kernels) and fp64 (generic kernels).
#include "msda_launch.cuh"

namespace msda {

MSDA_DECLARE_TYPED(, float, float)

int forward_f64(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                const void* attn, void* out, const msda_dims* d, cudaStream_t st)
{
    return launch_forward_generic<double, double, double>(value, shapes, ls, loc, attn, out, d, st);
}

int backward_f64(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                 const void* attn, const void* go, void* gv, void* gl, void* ga,
                 const msda_dims* d, cudaStream_t st)
{
    const size_t nval = (size_t)d->N * d->S * d->M * d->D;
    cudaError_t e = cudaMemsetAsync(gv, 0, nval * sizeof(double), st);
    if (e != cudaSuccess) return (int)e;
    if (d->Lq == 0) return 0;
    return launch_backward_generic<double, double, double>(value, shapes, ls, loc, attn, go,
                                                           (double*)gv, gl, ga, d, st);
}

}  // namespace msda
