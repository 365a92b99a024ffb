// msda_static.cu -- msda_b200_backward_static_scratch: the 16-bit backward with its scratch memory (fp32 grad_value
// accumulator, the dense-window kernel's work counter and done-item flags) in a device array that is part of
// the library image instead of a caller-provided workspace.  No counterpart in the reference; see
// include/msda_b200.h.  For callers that cannot allocate (or cannot size a workspace ahead of time): the call
// itself allocates nothing, and there is nothing for the caller to allocate either.
//
// The array's name is a HOST-side stand-in in host code, not a device address: the launchers must be given the
// device address cudaGetSymbolAddress returns, resolved once per device with that device current.
#include <mutex>

#include "msda_launch.cuh"

namespace msda {

// instantiated in msda_inst_{f32,bf16,f16}.cu
MSDA_DECLARE_TYPED(extern, float, float)
MSDA_DECLARE_TYPED(extern, __nv_bfloat16, float)
MSDA_DECLARE_TYPED(extern, __nv_bfloat16, __nv_bfloat16)
MSDA_DECLARE_TYPED(extern, __half, float)
MSDA_DECLARE_TYPED(extern, __half, __half)

// Held by every GPU that loads this module until the process exits.
static __device__ __align__(256) unsigned char g_static_scratch[MSDA_B200_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static void* g_scratch_addr[kMaxDevices];       // device address of g_static_scratch on device i (0 = not resolved)
static std::mutex g_scratch_mutex;

// Device address of the scratch array on the CURRENT device.
static int static_scratch(void** out)
{
    int dev = 0;
    if (cudaError_t e = cudaGetDevice(&dev)) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        if (cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch)) return (int)e;
        if (!p) return (int)cudaErrorInvalidSymbol;
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace msda

using namespace msda;

extern "C" int msda_b200_backward_static_scratch(const void* value, int value_dtype, const int64_t* spatial_shapes,
                                                 const int64_t* level_start, const void* sampling_loc,
                                                 const void* attn_weight, int aux_dtype, const void* grad_output,
                                                 void* grad_value, void* grad_sampling_loc, void* grad_attn_weight,
                                                 const msda_dims* dims, void* stream)
{
    g_last_launches = 0;
    if (int rc = check_common(value, spatial_shapes, sampling_loc, attn_weight, dims)) return rc;
    if (!grad_value) return MSDA_E_NULL;
    if (dims->Lq > 0 && (!grad_output || !grad_sampling_loc || !grad_attn_weight)) return MSDA_E_NULL;
    const size_t need = msda_b200_backward_workspace_bytes(value_dtype, dims);
    if (need > (size_t)MSDA_B200_STATIC_SCRATCH_BYTES) return MSDA_E_SCRATCH_LIMIT;
    void* ws = nullptr;
    if (need > 0)
        if (int rc = static_scratch(&ws)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // backward_typed zero-fills the part of the scratch it uses (accumulator, counter, flags) on `st` before the
    // kernels: what the previous call left behind is never read.
#define MSDA_BWD(VT, AT) \
    return backward_typed<VT, AT>(value, spatial_shapes, level_start, sampling_loc, attn_weight, grad_output, \
                                  grad_value, grad_sampling_loc, grad_attn_weight, ws, need, dims, st, nullptr)
    if (value_dtype == MSDA_F32 && aux_dtype == MSDA_F32) MSDA_BWD(float, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_F32) MSDA_BWD(__nv_bfloat16, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_BF16) MSDA_BWD(__nv_bfloat16, __nv_bfloat16);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F32) MSDA_BWD(__half, float);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F16) MSDA_BWD(__half, __half);
#undef MSDA_BWD
    return MSDA_E_DTYPE;
}
