"""MSDeformAttn -- the operator module, mirroring the reference's
src/trackformer/models/ops/modules/ms_deform_attn.py:15-89 (same constructor, parameter
names/shapes/initialisation, forward signature incl. ``query_attn_mask``), so TrackFormer /
KineTTracker checkpoints load unchanged and deformable_transformer.py can import it as is.

The four Linear layers stay cuBLAS GEMMs (their bias gradients are fixed-order column sums,
functions/layer_ops.py); the sampling + weighted reduction runs in the
sm_100a kernels through MSDeformAttnFunction.  Reference quirk kept on purpose: 2-d reference
points are normalised by ``input_spatial_shapes`` in (H, W) order (:77-79), i.e. the x offset is
divided by H and the y offset by W -- upstream Deformable-DETR flips to (W, H); the reference
does not, and parity is defined against the reference.
"""

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn.init import constant_, xavier_uniform_

from .._cache import PerTensorCache
from ..functions import MSDeformAttnFunction, MSDeformAttnFusedFunction, fused_supported
from ..functions.layer_ops import invalidate_lp_cache, linear as _linear


_shapes_validated = PerTensorCache()   # spatial_shapes tensor object -> True, keyed additionally by Len_in


def _check_spatial_shapes(shapes, len_in):
    """The reference's ``assert (shapes[:, 0] * shapes[:, 1]).sum() == Len_in`` (ms_deform_attn.py:62).
    On CUDA tensors the comparison is a device->host sync per call, 12 per frame in a 6+6-layer
    transformer although every layer receives the same tensor.  The result is remembered for the tensor
    objects it was computed on (identity + version counter, shared by all module instances; a few of them, so
    the multi-frame configuration's alternation between the encoder's and the decoder's level list hits too), so
    a frame pays for it once; while a CUDA graph is being captured a sync is illegal and the check is skipped
    (capture needs an eager warm-up run with the same tensors anyway, which validates them)."""
    if _shapes_validated.get(shapes, len_in):
        return
    if shapes.is_cuda and torch.cuda.is_current_stream_capturing():
        return
    assert (shapes[:, 0] * shapes[:, 1]).sum() == len_in
    _shapes_validated.put(shapes, True, len_in)


_mask_empty = PerTensorCache()        # padding-mask tensor object -> True when it masks nothing


def _no_padding(mask):
    """True when the (N, S) padding mask is all False -- tracking inference and any batch of equally sized images.
    The reference zero-fills ``value`` through the mask in every layer regardless (ms_deform_attn.py:65-66): a
    read and a write of the whole value tensor, 12 times per frame, for nothing.  The answer is one device->host
    read per mask tensor OBJECT (every layer of a forward receives the same one); during CUDA-graph capture no
    read is possible and an unknown mask is applied."""
    hit = _mask_empty.get(mask)
    if hit is not None:
        return hit
    if mask.is_cuda and torch.cuda.is_current_stream_capturing():
        return False
    return _mask_empty.put(mask, not bool(mask.any()))


class MSDeformAttn(nn.Module):
    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4, im2col_step=64):
        super().__init__()
        assert d_model % n_heads == 0
        # True: softmax, query mask and the sampling-location arithmetic below run inside the CUDA
        # kernels (MSDeformAttnFusedFunction) whenever the library has a fused kernel for the shape;
        # False: the reference's own sequence of PyTorch ops followed by MSDeformAttnFunction.
        self.fused_prologue = True

        self.im2col_step = im2col_step
        self.d_model = d_model
        self.n_levels = n_levels
        self.n_heads = n_heads
        self.n_points = n_points

        self.sampling_offsets = nn.Linear(d_model, n_heads * n_levels * n_points * 2)
        self.attention_weights = nn.Linear(d_model, n_heads * n_levels * n_points)
        self.value_proj = nn.Linear(d_model, d_model)
        self.output_proj = nn.Linear(d_model, d_model)

        self._reset_parameters()

    def _reset_parameters(self):
        # reference :34-47: zero offset weights; bias = the 8 compass directions cycled over the
        # heads, scaled by (point index + 1); zero attention logits; xavier projections.
        constant_(self.sampling_offsets.weight.data, 0.)
        compass = torch.tensor([[-1., -1.], [-1., 0.], [-1., 1.], [0., -1.],
                                [0., 1.], [1., -1.], [1., 0.], [1., 1.]])
        # the reference's .view(n_heads, 1, 1, 2) of a 16-element tensor fixes n_heads == 8
        assert self.n_heads == 8, "the reference initialisation (ms_deform_attn.py:36-37) requires n_heads == 8"
        grid_init = compass.view(self.n_heads, 1, 1, 2).repeat(1, self.n_levels, self.n_points, 1)
        scale = torch.arange(1, self.n_points + 1, dtype=torch.float32).view(1, 1, self.n_points, 1)
        grid_init = grid_init * scale
        with torch.no_grad():
            self.sampling_offsets.bias = nn.Parameter(grid_init.reshape(-1))
        constant_(self.attention_weights.weight.data, 0.)
        constant_(self.attention_weights.bias.data, 0.)
        xavier_uniform_(self.value_proj.weight.data)
        constant_(self.value_proj.bias.data, 0.)
        xavier_uniform_(self.output_proj.weight.data)
        constant_(self.output_proj.bias.data, 0.)
        invalidate_lp_cache(self)       # the .data edits above are invisible to the low-precision copies' cache key

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes,
                input_padding_mask=None, query_attn_mask=None):
        """
        query (N, Lq, C); reference_points (N, Lq, L, 2) in [0,1] or (N, Lq, L, 4) boxes;
        input_flatten (N, S, C); input_spatial_shapes (L, 2) [(H_l, W_l)];
        input_padding_mask (N, S) True = padding; query_attn_mask (N, Lq) True = masked query.
        -> (N, Lq, C)
        """
        N, Len_q, _ = query.shape
        N, Len_in, _ = input_flatten.shape
        _check_spatial_shapes(input_spatial_shapes, Len_in)

        value = _linear(input_flatten, self.value_proj)
        if input_padding_mask is not None and not _no_padding(input_padding_mask):
            value = value.masked_fill(input_padding_mask[..., None], float(0))
        value = value.view(N, Len_in, self.n_heads, self.d_model // self.n_heads)

        sampling_offsets = _linear(query, self.sampling_offsets).view(
            N, Len_q, self.n_heads, self.n_levels, self.n_points, 2)
        attention_weights = _linear(query, self.attention_weights).view(
            N, Len_q, self.n_heads, self.n_levels * self.n_points)
        if reference_points.shape[-1] not in (2, 4):
            raise ValueError(
                'Last dim of reference_points must be 2 or 4, but get {} instead.'.format(
                    reference_points.shape[-1]))
        if self.fused_prologue and fused_supported(value, self.n_levels, self.n_points):
            # same arithmetic as the lines below (reference :69-82), done in the kernels' owner lanes
            output = MSDeformAttnFusedFunction.apply(
                value, input_spatial_shapes, reference_points, sampling_offsets, attention_weights,
                query_attn_mask, input_spatial_shapes, self.im2col_step)
            return _linear(output, self.output_proj, skip_bias_grad=getattr(self, "_kinet_skip_output_bias_grad", False))
        attention_weights = F.softmax(attention_weights, -1).view(
            N, Len_q, self.n_heads, self.n_levels, self.n_points)

        if query_attn_mask is not None:
            attention_weights = attention_weights.masked_fill(
                query_attn_mask[..., None, None, None], float(0))

        if reference_points.shape[-1] == 2:
            sampling_locations = reference_points[:, :, None, :, None, :] \
                + sampling_offsets / input_spatial_shapes[None, None, None, :, None, :]
        elif reference_points.shape[-1] == 4:
            sampling_locations = reference_points[:, :, None, :, None, :2] \
                + sampling_offsets / self.n_points * reference_points[:, :, None, :, None, 2:] * 0.5
        else:
            raise ValueError(
                'Last dim of reference_points must be 2 or 4, but get {} instead.'.format(
                    reference_points.shape[-1]))
        output = MSDeformAttnFunction.apply(
            value, input_spatial_shapes, sampling_locations, attention_weights, self.im2col_step)
        return _linear(output, self.output_proj, skip_bias_grad=getattr(self, "_kinet_skip_output_bias_grad", False))
