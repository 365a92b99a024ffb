/*
 * kinet_layers_b200.h -- C ABI of the glue kernels of the deformable-transformer layers that CALL
 * the MSDeformAttn op.  Same library as msda_b200.h
 * (libmsda_b200.so), same conventions: device pointers, plain sizes, work is enqueued on `stream`
 * (cudaStream_t as void*), nothing is allocated, nothing synchronises, nothing throws; return 0,
 * a negative MSDA_E_* or a positive cudaError_t.
 *
 * What they replace in the reference (LuisCossioUOH/Kinet, src/trackformer/models/
 * deformable_transformer.py), per residual branch of a layer:
 *
 *   src = src + dropout(src2); src = norm(src)        :294-296, :282-284 (encoder), :375-376, :382-383,
 *                                                     :363-365 (decoder)
 *   with_pos_embed(src, pos)                          :277-278, :293, :369, :379   (input of the next op)
 *   + the dtype casts autocast inserts in front of the next Linear layers
 *
 * i.e. an fp32 add, a LayerNorm, another fp32 add and up to three cast kernels (and their autograd
 * twins, of which LayerNorm's gamma/beta reduction and the Linear bias reductions are the slowest
 * kernels of a training step) become ONE pass forward and ONE pass backward over the rows.
 *
 * Element types use the MSDA_F32 / MSDA_BF16 / MSDA_F16 codes of msda_b200.h (no fp64).
 */
#ifndef KINET_LAYERS_B200_H
#define KINET_LAYERS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KINET_LN_MAX_C 1024 /* channels per row: multiple of 4, <= 1024 */

/* y = LayerNorm(x + r) * gamma + beta ; y_lp = cast(y) ; ypos_lp = cast(y + pos).
 * Rows are [rows][C] contiguous.  Statistics and the affine transform are computed in fp32 on
 * the fp32 sum x + r (two-pass variance over registers), as torch.nn.functional.layer_norm does. */
typedef struct kinet_ln_fwd_args {
    const void* x;      int32_t x_dtype;   /* residual stream in                          */
    const void* r;      int32_t r_dtype;   /* branch output added to x; NULL = none        */
    const void* pos;    int32_t pos_dtype; /* positional embedding; NULL = no ypos_lp      */
    int64_t pos_rows;                      /* row i uses pos[i % pos_rows] (broadcast over the batch) */
    const float* gamma; const float* beta; /* [C] fp32                                    */
    void* y;            int32_t y_dtype;   /* LayerNorm output (residual stream out)      */
    void* y_lp;                            /* optional copy of y in lp_dtype               */
    void* ypos_lp;      int32_t lp_dtype;  /* optional y + pos in lp_dtype                 */
    float* mean_rstd;                      /* optional [rows][2] saved for the backward    */
    int64_t rows;       int32_t C;
    float eps;
    /* ABI 3: dropout of the branch inside the kernel, y = LayerNorm(x + dropout(r)) -- the reference's
     * src = src + self.dropoutN(src2); src = self.normN(src) (deformable_transformer.py:294-296 ...) without a
     * dropout kernel and its mask tensor.  drop_p = 0: none.  An element is kept with probability 1 - p' (p' = p
     * rounded to 1 / 65536) and scaled by 1 / (1 - p'); which elements are kept is a hash of (element index,
     * drop_seed, drop_site, *drop_step): the backward, given the same four values, regenerates the mask.
     * drop_step may be NULL (= 0); as a DEVICE pointer it lets a replayed CUDA graph draw a new mask every step. */
    float drop_p;       uint32_t drop_site;
    uint64_t drop_seed; const int64_t* drop_step;
} kinet_ln_fwd_args;

int kinet_b200_add_layernorm_forward(const kinet_ln_fwd_args* args, void* stream);

/* Backward of the above.  The incoming gradient of a row is dy + dy_lp + dypos_lp (each optional).
 * dx (x_dtype) and dr (r_dtype, optional) receive the same values d(x + r); dgamma_dbeta [2][C]
 * fp32 is fully overwritten.  `partial` is a caller-provided fp32 scratch of
 * kinet_b200_colsum_partial_floats(2 * C) elements (per-CTA column sums, reduced in a fixed order:
 * the result does not depend on scheduling). */
typedef struct kinet_ln_bwd_args {
    const void* dy;       int32_t dy_dtype;
    const void* dy_lp;    const void* dypos_lp; int32_t lp_dtype;
    const void* x;        int32_t x_dtype;
    const void* r;        int32_t r_dtype;
    const float* gamma;
    const float* mean_rstd;
    void* dx;
    void* dr;
    float* dgamma_dbeta;
    float* partial;
    int64_t rows;         int32_t C;
    /* ABI 3: the forward's dropout arguments (dr = mask * d(x + r) / (1 - p')), and want_dbias: dgamma_dbeta is
     * [3][C] and its third row receives the column sums of dr -- the bias gradient of the Linear layer that produced
     * the branch (output_proj, linear2), which then needs no reduction pass of its own; `partial` must hold
     * kinet_b200_colsum_partial_floats(3 * C) elements in that case. */
    int32_t want_dbias;
    float drop_p;         uint32_t drop_site;
    uint64_t drop_seed;   const int64_t* drop_step;
} kinet_ln_bwd_args;

int kinet_b200_add_layernorm_backward(const kinet_ln_bwd_args* args, void* stream);

/* Column sums out[c] = sum_i a[i][c] of a [rows][C] matrix (the bias gradient of a Linear layer,
 * which PyTorch computes with a generic reduce kernel): fp32 accumulation, fixed summation order.
 * C: multiple of 4, <= 1024.  out is out_dtype [C]; partial as above with n = C (not touched when rows <= 2048:
 * few rows are summed by one launch without scratch). */
size_t kinet_b200_colsum_partial_floats(int32_t n_columns);
int kinet_b200_colsum(const void* a, int32_t a_dtype, int64_t rows, int32_t C,
                      void* out, int32_t out_dtype, float* partial, void* stream);

/* ReLU backward fused with the column sum: dh = dy * (h > 0) is written (dtype of dy / h) and summed
 * over the rows in the same pass -- the bias gradient of the Linear layer feeding a ReLU
 * (deformable_transformer.py:283, :362: linear2(dropout(activation(linear1(x))))).  h is the ReLU's OUTPUT. */
int kinet_b200_relu_backward_colsum(const void* dy, const void* h, int32_t dtype, int64_t rows, int32_t C,
                                    void* dh, void* out, int32_t out_dtype, float* partial, void* stream);

/* Dropout without a mask tensor (ABI 3): out = dropout(x, drop_p) elementwise over n elements (n % 4 == 0), the keep / drop
 * decision hashed from (element index, drop_seed, drop_site, *drop_step) exactly as in the add + LayerNorm kernels.  Meant for
 * the FFN's inner dropout, linear2(dropout(activation(linear1(x)))) (deformable_transformer.py:283, :362): for a ReLU the
 * backward needs neither the mask nor the hash, because dropout(relu(a)) is positive exactly where the element survived and
 * the ReLU was open -- kinet_b200_relu_dropout_backward_colsum(dy, h_dropped, drop_p, ...) writes
 * dh = dy * (h_dropped > 0) / (1 - p') and its column sums in one pass (in place of PyTorch's masked_scale kernel, the ReLU
 * backward and the bias-gradient reduction). */
int kinet_b200_dropout_forward(const void* x, void* out, int32_t dtype, int64_t n, float drop_p, uint32_t drop_site,
                               uint64_t drop_seed, const int64_t* drop_step, void* stream);
int kinet_b200_relu_dropout_backward_colsum(const void* dy, const void* h_dropped, float drop_p, int32_t dtype, int64_t rows,
                                            int32_t C, void* dh, void* out, int32_t out_dtype, float* partial, void* stream);

/* ---- attention core of nn.MultiheadAttention for short sequences: the `kine` path ----
 *
 * Replaces, inside the reference's KinematicTransformer / DualKinematicTransformer
 * (src/trackformer/models/transformer.py:85-185), what nn.MultiheadAttention does between its input and output
 * projections in TransformerEncoderLayer.forward_post (:331-333) and TransformerDecoderLayer.forward_post
 * (:393-401): out[n][i][h][:] = sum_j softmax_j(scale * <q[n][i][h], k[n][j][h]> + (-inf if key j is masked)) v[n][j][h][:]
 * for <= 150 object + T tracklet tokens, 8 heads of 36 channels (cfgs/train_kinet.yaml:17-18).  The projections stay
 * cuBLAS GEMMs.  fp32 arithmetic for every storage type; D <= KINET_MHA_MAX_D; any Lq, Lk (keys are walked in tiles
 * with an online softmax).  A query whose keys are ALL masked yields 0 (torch >= 2.5; torch 1.5 returned NaN).
 *
 * Element (n, i, h, c) of q lives at q + n * q_batch_stride + i * q_row_stride + h * D + c (elements): batch-first
 * and sequence-first layouts, and slices of a packed q|k|v GEMM output, are all expressible.  Same for k, v, out. */
#define KINET_MHA_MAX_D 64

typedef struct kinet_mha_args {
    const void* q; const void* k; const void* v;
    int64_t q_batch_stride, q_row_stride, k_batch_stride, k_row_stride, v_batch_stride, v_row_stride;
    const uint8_t* key_padding_mask;   /* [N][Lk], nonzero = key masked; NULL = none                              */
    void* out;                         /* element (n, i, h, c) at out + n * out_batch_stride + i * out_row_stride + h * D + c */
    int64_t out_batch_stride, out_row_stride;
    float* lse;                        /* optional [N][H][Lq] fp32: log-sum-exp of the scaled logits (the backward reads it) */
    int32_t dtype;                     /* MSDA_F32 / MSDA_BF16 / MSDA_F16: q, k, v, out (and the backward's tensors)  */
    int32_t N, H, D, Lq, Lk;
    float scale;                       /* 1 / sqrt(D) for nn.MultiheadAttention                                   */
} kinet_mha_args;

int kinet_b200_small_mha_forward(const kinet_mha_args* args, void* stream);

/* Backward: fwd holds the forward's arguments (out and lse as the forward wrote them).  dq [N][Lq][H*D], dk and dv
 * [N][Lk][H*D] are contiguous, element type fwd.dtype, fully overwritten; two kernels (dK / dV per key tile, dQ per
 * query tile), probabilities recomputed from lse, one writer per element (no atomics: bit-reproducible).
 * Rows whose keys were all masked get zero gradients. */
typedef struct kinet_mha_bwd_args {
    kinet_mha_args fwd;
    const void* dout;                  /* gradient of out, addressed like out with its own strides                 */
    int64_t dout_batch_stride, dout_row_stride;
    void* dq; void* dk; void* dv;
} kinet_mha_bwd_args;

int kinet_b200_small_mha_backward(const kinet_mha_bwd_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif
