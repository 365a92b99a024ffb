/*
 * msda_b200.h -- C ABI of the B200-native multi-scale deformable attention op.
 *
 * This is the drop-in boundary for the reference's native extension
 * `MultiScaleDeformableAttention` (LuisCossioUOH/Kinet,
 * src/trackformer/models/ops/src/).  Entry points replace:
 *
 *   msda_b200_forward   <- ms_deform_attn_forward   (src/vision.cpp:5,
 *                          src/ms_deform_attn.h:10-28,
 *                          src/cuda/ms_deform_attn_cuda.cu:19-86)
 *   msda_b200_backward  <- ms_deform_attn_backward  (src/vision.cpp:6,
 *                          src/ms_deform_attn.h:30-49,
 *                          src/cuda/ms_deform_attn_cuda.cu:89-168)
 *
 * Conventions: plain pointers and sizes only; all data pointers are DEVICE
 * pointers on the current CUDA device; calls enqueue work on `stream`
 * (a cudaStream_t passed as void*; NULL = legacy default stream) and return
 * without synchronising.  Nothing is allocated; nothing throws.  Return value
 * is 0 on success, a negative MSDA_E_* for argument errors, or a positive
 * cudaError_t from the launch.  There is no CPU implementation: on a machine
 * without a CUDA device every compute entry point returns a CUDA error.
 *
 * Tensor layouts (row-major, contiguous; same as the reference,
 * ms_deform_attn_cuda.cu:25-27):
 *   value            [N][S][M][D]          value_dtype
 *   spatial_shapes   [L][2] int64 (H, W)   device memory
 *   level_start      [L] int64 or NULL     device memory; NULL = computed from
 *                                          spatial_shapes inside the kernel
 *                                          (the reference recomputes it on the
 *                                          host every call, :52-58)
 *   sampling_loc     [N][Lq][M][L][P][2]   aux_dtype, (x, y) in [0,1]
 *   attn_weight      [N][Lq][M][L][P]      aux_dtype
 *   output           [N][Lq][M*D]          value_dtype
 *   grad_output      [N][Lq][M*D]          value_dtype
 *   grad_value       like value            value_dtype  (fully overwritten)
 *   grad_sampling_loc, grad_attn_weight    aux_dtype    (fully overwritten)
 */
#ifndef MSDA_B200_H
#define MSDA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSDA_B200_ABI_VERSION 3

/* element types */
#define MSDA_F32 0
#define MSDA_BF16 1
#define MSDA_F64 2 /* parity with the reference's AT_DISPATCH_FLOATING_TYPES (gradcheck) */
#define MSDA_F16 3

/* argument errors (negative so they never collide with cudaError_t) */
#define MSDA_E_DTYPE (-1)   /* unsupported value_dtype / aux_dtype combination */
#define MSDA_E_SHAPE (-2)   /* non-positive or unsupported dimension */
#define MSDA_E_NULL (-3)    /* required pointer is NULL */
#define MSDA_E_ALIGN (-4)   /* pointer not 16-byte aligned */
#define MSDA_E_WORKSPACE (-5) /* workspace too small */
#define MSDA_E_SCRATCH_LIMIT (-6) /* msda_b200_backward_static_scratch: the call needs more than MSDA_B200_STATIC_SCRATCH_BYTES */

#define MSDA_MAX_LEVELS 16

/* Supported (value_dtype, aux_dtype): (F32,F32) (BF16,F32) (BF16,BF16)
 * (F16,F32) (F16,F16) (F64,F64).  D*sizeof(value) must be a multiple of 16
 * for the vector paths (D % 4 == 0 fp32, D % 8 == 0 bf16/f16); any other D
 * takes the scalar path.  1 <= L <= MSDA_MAX_LEVELS. */

typedef struct msda_dims {
    int32_t N;  /* batch */
    int32_t S;  /* sum_l H_l*W_l */
    int32_t M;  /* heads */
    int32_t D;  /* channels per head */
    int32_t L;  /* levels */
    int32_t Lq; /* queries (0 allowed: no-op for outputs, grad_value zero-filled) */
    int32_t P;  /* points per level */
} msda_dims;

int msda_b200_abi_version(void);

/* Human-readable text for a return code of this library. */
const char *msda_b200_error_string(int code);

/* output = sum_{l,p} attn * bilinear(value_l, loc)   (fused, no scratch) */
int msda_b200_forward(const void *value, int value_dtype,
                      const int64_t *spatial_shapes, const int64_t *level_start,
                      const void *sampling_loc, const void *attn_weight, int aux_dtype,
                      void *output, const msda_dims *dims, void *stream);

/* Bytes of device scratch msda_b200_backward needs (0 for f32/f64 value:
 * gradients accumulate in place; bf16/f16 accumulate grad_value in fp32). */
size_t msda_b200_backward_workspace_bytes(int value_dtype, const msda_dims *dims);

/* All three gradients in one call.  grad_value/grad_sampling_loc/grad_attn_weight
 * need not be initialised.  workspace may be NULL when the size above is 0. */
int msda_b200_backward(const void *value, int value_dtype,
                       const int64_t *spatial_shapes, const int64_t *level_start,
                       const void *sampling_loc, const void *attn_weight, int aux_dtype,
                       const void *grad_output,
                       void *grad_value, void *grad_sampling_loc, void *grad_attn_weight,
                       void *workspace, size_t workspace_bytes,
                       const msda_dims *dims, void *stream);

/* ---- Backward without a workspace and without any allocation ---------------------------------------
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, results and return codes as msda_b200_backward minus `workspace` / `workspace_bytes`:
 * the scratch the 16-bit backward needs (fp32 grad_value accumulator + the dense-window kernel's counter
 * and flags, i.e. msda_b200_backward_workspace_bytes) lives in ONE file-scope `static __device__` array of
 * MSDA_B200_STATIC_SCRATCH_BYTES inside the library image; every GPU that loads the library holds it
 * until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library
 * routine); the array's device address is resolved once per device (cudaGetSymbolAddress, with that
 * device current) on the first call.
 *   - a call that needs more scratch than the array holds returns MSDA_E_SCRATCH_LIMIT and launches
 *     nothing (32 MiB: N*S*M*D <= ~8.3 M value elements, e.g. the tracking configuration N=1, S=20906,
 *     M=8, D=36; the 800x1333 training batch N=2, D=32 needs 91 MB -> use msda_b200_backward);
 *   - every call on a device shares the array and finds the previous call's contents in it; the part a
 *     call uses is zero-filled on `stream` before its kernels, so calls that are ordered on the GPU
 *     (same stream, or streams ordered by events) are independent of each other;
 *   - calls that OVERLAP on the GPU (different streams, no ordering between them; concurrent graph
 *     launches; several host threads with their own streams) race on the array and produce garbage:
 *     serialise them or use msda_b200_backward with one workspace per stream;
 *   - f32 values need no scratch (the call equals msda_b200_backward with a NULL workspace); f64 is not
 *     offered here (MSDA_E_DTYPE). */
#define MSDA_B200_STATIC_SCRATCH_BYTES (32u * 1024u * 1024u)

int msda_b200_backward_static_scratch(const void *value, int value_dtype,
                                      const int64_t *spatial_shapes, const int64_t *level_start,
                                      const void *sampling_loc, const void *attn_weight, int aux_dtype,
                                      const void *grad_output,
                                      void *grad_value, void *grad_sampling_loc, void *grad_attn_weight,
                                      const msda_dims *dims, void *stream);

/* ---- Fused module prologue -------------------------------------------------
 * The reference's MSDeformAttn.forward (ops/modules/ms_deform_attn.py:69-82) materialises softmax'ed
 * attention weights and sampling locations with ~6 elementwise kernels before it calls the op.  The
 * fused entry points read what the module's Linear layers produce instead and do that arithmetic in
 * the kernel:
 *   attention[n,q,m,:]  = softmax over the L*P logits of (n,q,m); 0 for queries with query_mask != 0
 *   location[n,q,m,l,p] = ref[n,q,l,:] + offset[n,q,m,l,p,:] / offset_norm[l,:]            (ref_dim 2)
 *                       = ref[n,q,l,:2] + offset / P * ref[n,q,l,2:] * 0.5                 (ref_dim 4)
 * offset_norm is what the module divides by: the reference passes input_spatial_shapes, i.e. (H_l, W_l)
 * for the (x, y) offsets (:77-79; upstream Deformable-DETR uses (W_l, H_l)) -- the caller decides.
 * The backward returns the gradients w.r.t. the raw offsets and logits (through the location
 * arithmetic and the softmax) and, when grad_sampling_loc != NULL, d out / d location in fp32
 * ([N][Lq][M][L][P][2]) from which the caller reduces the reference-point gradient.
 * Return MSDA_E_SHAPE = no fused kernel for this shape (D not in {16,32,36,64}, misaligned pointers, or
 * L*P > 4 * 2^floor(log2(D/4)) -- D/8 for the 16-bit forward): run the prologue outside and call the plain
 * entry points. */
typedef struct msda_fused_inputs {
    const void *sampling_offsets;     /* [N][Lq][M][L][P][2]  aux_dtype, raw Linear output */
    const void *attn_logits;          /* [N][Lq][M][L*P]      aux_dtype, before softmax */
    const float *reference_points;    /* [N][Lq][L][ref_dim]  fp32, 16-byte aligned */
    const float *offset_norm;         /* [L][2] fp32 divisors of (x, y) offsets; ref_dim == 2 only */
    const unsigned char *query_mask;  /* [N][Lq] or NULL; nonzero = attention weights of the query zeroed */
    int32_t ref_dim;                  /* 2 or 4 */
} msda_fused_inputs;

int msda_b200_forward_fused(const void *value, int value_dtype,
                            const int64_t *spatial_shapes, const int64_t *level_start,
                            const msda_fused_inputs *inputs, int aux_dtype,
                            void *output, const msda_dims *dims, void *stream);

int msda_b200_backward_fused(const void *value, int value_dtype,
                             const int64_t *spatial_shapes, const int64_t *level_start,
                             const msda_fused_inputs *inputs, int aux_dtype,
                             const void *grad_output,
                             void *grad_value, void *grad_offsets, void *grad_logits,
                             float *grad_sampling_loc /* optional */,
                             void *workspace, size_t workspace_bytes,
                             const msda_dims *dims, void *stream);

/* Number of kernels the last forward / backward call of this thread launched
 * (bench.py's gpu_launches evidence). */
int msda_b200_last_launch_count(void);

/* Measurement hook: the NEXT forward or backward call of this thread records the two
 * cudaEvent_t handles (passed as void*) on its stream immediately before and after its
 * dominant kernel (the gather kernel / the gather+scatter kernel), so a caller can time that
 * kernel alone inside a larger timed region.  One-shot; pass NULLs to cancel. */
void msda_b200_trace_next(void *start_event, void *stop_event);

#ifdef __cplusplus
}
#endif
#endif /* MSDA_B200_H */
