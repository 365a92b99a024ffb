// msda_abi.cu -- extern "C" entry points of libmsda_b200.so (see include/msda_b200.h).
//
// Host side of the reference's ms_deform_attn_cuda_forward / _backward
// (src/trackformer/models/ops/src/cuda/ms_deform_attn_cuda.cu:19-86, :89-168),
// re-designed: no allocation, no host loop over im2col_step chunks, no per-call
// ATen ops for level_start_index, no host synchronisation.  Picks a kernel
// instantiation from (value dtype, aux dtype, D) and a tiling from the shape.
#include "msda_launch.cuh"

namespace msda {

thread_local int g_last_launches = 0;
thread_local cudaEvent_t g_trace_start = nullptr, g_trace_stop = nullptr;

// instantiated in msda_inst_{f32,bf16,f16}.cu
MSDA_DECLARE_TYPED(extern, float, float)
MSDA_DECLARE_TYPED(extern, __nv_bfloat16, float)
MSDA_DECLARE_TYPED(extern, __nv_bfloat16, __nv_bfloat16)
MSDA_DECLARE_TYPED(extern, __half, float)
MSDA_DECLARE_TYPED(extern, __half, __half)

}  // namespace msda

using namespace msda;

extern "C" {

int msda_b200_abi_version(void) { return MSDA_B200_ABI_VERSION; }

const char* msda_b200_error_string(int code)
{
    switch (code) {
        case 0: return "success";
        case MSDA_E_DTYPE: return "msda_b200: unsupported value/aux dtype combination";
        case MSDA_E_SHAPE: return "msda_b200: invalid or unsupported dimension";
        case MSDA_E_NULL: return "msda_b200: required pointer is NULL";
        case MSDA_E_ALIGN: return "msda_b200: pointer is not 16-byte aligned";
        case MSDA_E_WORKSPACE: return "msda_b200: workspace missing or too small";
        case MSDA_E_SCRATCH_LIMIT:
            return "msda_b200: the call needs more scratch than the static array of msda_b200_backward_static_scratch "
                   "holds (MSDA_B200_STATIC_SCRATCH_BYTES); use msda_b200_backward with a workspace";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "msda_b200: unknown error";
}

int msda_b200_last_launch_count(void) { return g_last_launches; }



void msda_b200_trace_next(void* start_event, void* stop_event)
{
    g_trace_start = (cudaEvent_t)start_event;
    g_trace_stop = (cudaEvent_t)stop_event;
}

int msda_b200_forward(const void* value, int value_dtype, const int64_t* spatial_shapes,
                      const int64_t* level_start, const void* sampling_loc, const void* attn_weight,
                      int aux_dtype, void* output, const msda_dims* dims, void* stream)
{
    g_last_launches = 0;
    if (int rc = check_common(value, spatial_shapes, sampling_loc, attn_weight, dims)) return rc;
    if (dims->Lq == 0) return 0;
    if (!output) return MSDA_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
#define MSDA_FWD(VT, AT) \
    return forward_typed<VT, AT>(value, spatial_shapes, level_start, sampling_loc, attn_weight, output, dims, st, nullptr)
    if (value_dtype == MSDA_F32 && aux_dtype == MSDA_F32) MSDA_FWD(float, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_F32) MSDA_FWD(__nv_bfloat16, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_BF16) MSDA_FWD(__nv_bfloat16, __nv_bfloat16);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F32) MSDA_FWD(__half, float);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F16) MSDA_FWD(__half, __half);
#undef MSDA_FWD
    if (value_dtype == MSDA_F64 && aux_dtype == MSDA_F64)
        return forward_f64(value, spatial_shapes, level_start, sampling_loc, attn_weight, output, dims, st);
    return MSDA_E_DTYPE;
}

size_t msda_b200_backward_workspace_bytes(int value_dtype, const msda_dims* d)
{
    if (!d) return 0;
    // fp32 accumulator + the dense-window kernel's work counter and done-item flags; nothing for sparse query
    // sets, which accumulate in the 16-bit output itself
    if (value_dtype == MSDA_BF16 || value_dtype == MSDA_F16) {
        if (sparse_16bit_regime(d)) return 0;
        return dense_counter_offset((size_t)d->N * d->S * d->M * d->D) + dense_extra_bytes(d);
    }
    return 0;
}

int msda_b200_backward(const void* value, int value_dtype, const int64_t* spatial_shapes,
                       const int64_t* level_start, const void* sampling_loc, const void* attn_weight,
                       int aux_dtype, const void* grad_output, void* grad_value,
                       void* grad_sampling_loc, void* grad_attn_weight, void* workspace,
                       size_t workspace_bytes, const msda_dims* dims, void* stream)
{
    g_last_launches = 0;
    if (int rc = check_common(value, spatial_shapes, sampling_loc, attn_weight, dims)) return rc;
    if (!grad_value) return MSDA_E_NULL;
    if (dims->Lq > 0 && (!grad_output || !grad_sampling_loc || !grad_attn_weight)) return MSDA_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
#define MSDA_BWD(VT, AT) \
    return backward_typed<VT, AT>(value, spatial_shapes, level_start, sampling_loc, attn_weight, grad_output, \
                                  grad_value, grad_sampling_loc, grad_attn_weight, workspace, workspace_bytes, dims, st, \
                                  nullptr)
    if (value_dtype == MSDA_F32 && aux_dtype == MSDA_F32) MSDA_BWD(float, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_F32) MSDA_BWD(__nv_bfloat16, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_BF16) MSDA_BWD(__nv_bfloat16, __nv_bfloat16);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F32) MSDA_BWD(__half, float);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F16) MSDA_BWD(__half, __half);
#undef MSDA_BWD
    if (value_dtype == MSDA_F64 && aux_dtype == MSDA_F64)
        return backward_f64(value, spatial_shapes, level_start, sampling_loc, attn_weight, grad_output,
                            grad_value, grad_sampling_loc, grad_attn_weight, dims, st);
    return MSDA_E_DTYPE;
}


// ---- fused prologue ------------------------------------------------------------------

static int check_fused(const msda_fused_inputs* in, const msda_dims* dims, FusedArgs* fa)
{
    if (!in) return MSDA_E_NULL;
    if (dims->Lq > 0 && (!in->sampling_offsets || !in->attn_logits || !in->reference_points)) return MSDA_E_NULL;
    if (in->ref_dim != 2 && in->ref_dim != 4) return MSDA_E_SHAPE;
    if (in->ref_dim == 2 && !in->offset_norm) return MSDA_E_NULL;
    if ((reinterpret_cast<uintptr_t>(in->reference_points) & 15u) || (reinterpret_cast<uintptr_t>(in->offset_norm) & 7u))
        return MSDA_E_ALIGN;
    fa->ref = in->reference_points;
    fa->norm = in->offset_norm;
    fa->qmask = in->query_mask;
    fa->grad_loc = nullptr;
    fa->ref_dim = in->ref_dim;
    return 0;
}

int msda_b200_forward_fused(const void* value, int value_dtype, const int64_t* spatial_shapes,
                            const int64_t* level_start, const msda_fused_inputs* in, int aux_dtype,
                            void* output, const msda_dims* dims, void* stream)
{
    g_last_launches = 0;
    if (!in) return MSDA_E_NULL;
    if (int rc = check_common(value, spatial_shapes, in->sampling_offsets, in->attn_logits, dims)) return rc;
    FusedArgs fa;
    if (int rc = check_fused(in, dims, &fa)) return rc;
    if (dims->Lq == 0) return 0;
    if (!output) return MSDA_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
#define MSDA_FWD(VT, AT) \
    return forward_typed<VT, AT>(value, spatial_shapes, level_start, in->sampling_offsets, in->attn_logits, output, dims, \
                                 st, &fa)
    if (value_dtype == MSDA_F32 && aux_dtype == MSDA_F32) MSDA_FWD(float, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_F32) MSDA_FWD(__nv_bfloat16, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_BF16) MSDA_FWD(__nv_bfloat16, __nv_bfloat16);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F32) MSDA_FWD(__half, float);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F16) MSDA_FWD(__half, __half);
#undef MSDA_FWD
    return MSDA_E_DTYPE;
}

int msda_b200_backward_fused(const void* value, int value_dtype, const int64_t* spatial_shapes,
                             const int64_t* level_start, const msda_fused_inputs* in, int aux_dtype,
                             const void* grad_output, void* grad_value, void* grad_offsets, void* grad_logits,
                             float* grad_sampling_loc, void* workspace, size_t workspace_bytes,
                             const msda_dims* dims, void* stream)
{
    g_last_launches = 0;
    if (!in) return MSDA_E_NULL;
    if (int rc = check_common(value, spatial_shapes, in->sampling_offsets, in->attn_logits, dims)) return rc;
    FusedArgs fa;
    if (int rc = check_fused(in, dims, &fa)) return rc;
    fa.grad_loc = grad_sampling_loc;
    if (!grad_value) return MSDA_E_NULL;
    if (dims->Lq > 0 && (!grad_output || !grad_offsets || !grad_logits)) return MSDA_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
#define MSDA_BWD(VT, AT) \
    return backward_typed<VT, AT>(value, spatial_shapes, level_start, in->sampling_offsets, in->attn_logits, grad_output, \
                                  grad_value, grad_offsets, grad_logits, workspace, workspace_bytes, dims, st, &fa)
    if (value_dtype == MSDA_F32 && aux_dtype == MSDA_F32) MSDA_BWD(float, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_F32) MSDA_BWD(__nv_bfloat16, float);
    if (value_dtype == MSDA_BF16 && aux_dtype == MSDA_BF16) MSDA_BWD(__nv_bfloat16, __nv_bfloat16);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F32) MSDA_BWD(__half, float);
    if (value_dtype == MSDA_F16 && aux_dtype == MSDA_F16) MSDA_BWD(__half, __half);
#undef MSDA_BWD
    return MSDA_E_DTYPE;
}

}  // extern "C"
