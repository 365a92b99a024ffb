"""Encoder / decoder layers of the deformable transformer -- the callers of MSDeformAttn -- mirroring
the reference's src/trackformer/models/deformable_transformer.py:

    DeformableTransformerEncoderLayer   :260-299      DeformableTransformerEncoder   :302-331
    DeformableTransformerDecoderLayer   :334-386      DeformableTransformerDecoder   :389-433

Same constructor arguments, sub-module / parameter names (``self_attn``, ``cross_attn``, ``norm1..3``,
``linear1/2``; checkpoints load unchanged), forward signatures and arithmetic.  What changes is how the
work between the GEMMs is executed on the GPU: every ``x = norm(x + dropout(branch))`` and the
``with_pos_embed`` that follows it run as one fused kernel (functions/layer_ops.py) that also emits,
under CUDA autocast, the low-precision copies the next Linear layers read -- so a layer no longer
launches separate add / LayerNorm / add / cast kernels (and, in training, LayerNorm's slow gamma/beta
reduction and the generic bias-gradient reductions).  The residual stream keeps its dtype (fp32 under
autocast, exactly as the reference's modules would behave under autocast); the GEMMs stay cuBLAS.

Inputs the fused kernels do not cover (CPU tensors, fp64 streams, C % 4 != 0 or C > 1024) raise from
the kernels' wrappers -- there is no PyTorch fallback behind these classes.
"""
import copy
import weakref

import torch
import torch.nn.functional as F
from torch import nn

from ..functions.layer_ops import add_layer_norm, inference_mirror, linear, linear_params
from .ms_deform_attn import MSDeformAttn


def _get_clones(module, n):
    return nn.ModuleList([copy.deepcopy(module) for _ in range(n)])


def _get_activation_fn(activation):
    """Same names and error as the reference (:563-571)."""
    if activation == "relu":
        return F.relu
    if activation == "gelu":
        return F.gelu
    if activation == "glu":
        return F.glu
    raise RuntimeError(F"activation should be relu/gelu, not {activation}.")


def inverse_sigmoid(x, eps=1e-5):
    """src/trackformer/util/misc.py:609-613."""
    x = x.clamp(min=0, max=1)
    return torch.log(x.clamp(min=eps) / (1 - x).clamp(min=eps))


def _lp_dtype(x):
    """dtype of the copies handed to the next GEMMs: the autocast dtype when CUDA autocast is on."""
    if x.is_cuda and torch.is_autocast_enabled("cuda"):
        return torch.get_autocast_dtype("cuda")
    return x.dtype


_grids = None   # (weakref to the shapes tensor, its version, device, per-level pixel-centre grids)


def _level_grids(spatial_shapes, device):
    """Flattened pixel-centre coordinates (y, x) of every level.  Reading H and W out of a CUDA
    ``spatial_shapes`` tensor is a device->host sync (the reference pays it on every forward, :311-314, and it
    is illegal while a CUDA graph is being captured), so the grids are remembered for the tensor object
    (identity + version counter) they were built from; an eager warm-up run with the same tensor fills them."""
    global _grids
    is_t = torch.is_tensor(spatial_shapes)
    if (is_t and _grids is not None and _grids[0]() is spatial_shapes and _grids[1] == spatial_shapes._version
            and _grids[2] == device):
        return _grids[3]
    out = []
    for H_, W_ in (spatial_shapes.tolist() if is_t else spatial_shapes):
        H_, W_ = int(H_), int(W_)
        ys = torch.linspace(0.5, H_ - 0.5, H_, dtype=torch.float32, device=device)
        xs = torch.linspace(0.5, W_ - 0.5, W_, dtype=torch.float32, device=device)
        gy, gx = torch.meshgrid(ys, xs, indexing="ij")
        out.append((gy.reshape(-1), gx.reshape(-1), H_, W_))
    if is_t:
        _grids = (weakref.ref(spatial_shapes), spatial_shapes._version, device, out)
    return out


def _act_linear(x, layer, activation):
    """activation(layer(x)); ReLU rides in the GEMM's epilogue (and, in the backward, in the bias-gradient pass)."""
    if activation is F.relu:
        return linear(x, layer, relu=True)
    return activation(linear(x, layer))


def _self_attention(mha, q_in, v_in, key_padding_mask):
    """nn.MultiheadAttention(q, k = q, v)[0] on batch-first (N, Lq, C) inputs, computed from the module's own
    parameters (in_proj_weight / in_proj_bias / out_proj: checkpoints unchanged): one GEMM for q and k, one for
    v, F.scaled_dot_product_attention, out_proj.  Same arithmetic as the reference's call
    (deformable_transformer.py:369-373, need_weights aside); the projections go through ``linear_params`` so their
    bias gradients are fused column sums instead of 17 us generic reductions over 600 rows."""
    N, Lq, C = q_in.shape
    H = mha.num_heads
    w, b = mha.in_proj_weight, mha.in_proj_bias
    mirror = inference_mirror(mha, w, b, q_in)
    m_qk = m_v = None
    if mirror is not None:
        m_qk, m_v = (mirror[0][:2 * C], mirror[1][:2 * C]), (mirror[0][2 * C:], mirror[1][2 * C:])
    qk = linear_params(q_in, w[:2 * C], b[:2 * C], False, m_qk)
    v = linear_params(v_in, w[2 * C:], b[2 * C:], False, m_v)
    q, k = qk.split(C, dim=-1)
    q, k, v = (t.reshape(N, Lq, H, C // H).transpose(1, 2) for t in (q, k, v))
    drop = mha.dropout if mha.training else 0.0
    d = C // H
    if key_padding_mask is None:
        o = F.scaled_dot_product_attention(q, k, v, dropout_p=drop)
    elif q.is_cuda and q.dtype in (torch.bfloat16, torch.float16):
        # An attn_mask takes 16-bit SDPA off the flash kernel (and with head_dim % 8 != 0, as D = 36, all the way
        # down to the fp32 math path: +0.35 ms per tracking frame).  A key-padding mask is a per-key additive bias,
        # so it rides in a spare head channel instead: q gets a constant 1 there, k gets -6e4 for masked keys (logit
        # <= -6e4 * scale <= -3750 for d <= 256: exp underflows to exactly 0) and 0 otherwise, v gets 0.  No mask,
        # flash kernel, same softmax.  (A row whose keys are ALL masked averages v instead of producing NaN.)
        dp = (d // 8 + 1) * 8
        qe, ke, ve = (F.pad(t, (0, dp - d)) for t in (q, k, v))
        qe[..., d] = 1.0
        ke[..., d] = (key_padding_mask.to(ke.dtype) * -6.0e4)[:, None, :]
        o = F.scaled_dot_product_attention(qe, ke, ve, dropout_p=drop, scale=d ** -0.5)[..., :d]
    else:
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=~key_padding_mask[:, None, None, :], dropout_p=drop)
    return linear(o.transpose(1, 2).reshape(N, Lq, C), mha.out_proj)


class _Stream:
    """Residual stream between fused steps: x (stream dtype), and the GEMM-side copies of x and x + pos."""
    __slots__ = ("x", "x_lp", "xpos_lp")

    def __init__(self, x, x_lp, xpos_lp):
        self.x, self.x_lp, self.xpos_lp = x, x_lp, xpos_lp

    @staticmethod
    def start(x, pos):
        lp = _lp_dtype(x)
        x_lp = x if lp == x.dtype else x.to(lp)
        if pos is None:
            return _Stream(x, x_lp, x_lp)
        xpos = x + pos
        return _Stream(x, x_lp, xpos if lp == xpos.dtype else xpos.to(lp))


def _norm_step(x, branch, norm, pos, want_pos):
    """x <- norm(x + branch); returns the next _Stream (copies only where the stream dtype differs)."""
    lp = _lp_dtype(x)
    same = lp == x.dtype
    y, y_lp, ypos = add_layer_norm(x, branch, norm, pos=pos if want_pos else None, lp_dtype=lp, want_lp=not same)
    if same:
        y_lp = y
    return _Stream(y, y_lp, ypos if ypos is not None else y_lp)


class DeformableTransformerEncoderLayer(nn.Module):
    def __init__(self, d_model=256, d_ffn=1024, dropout=0.1, activation="relu",
                 n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        # self attention
        self.self_attn = MSDeformAttn(d_model, n_levels, n_heads, n_points)
        self.dropout1 = nn.Dropout(dropout)
        self.norm1 = nn.LayerNorm(d_model)
        # ffn
        self.linear1 = nn.Linear(d_model, d_ffn)
        self.activation = _get_activation_fn(activation)
        self.dropout2 = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_ffn, d_model)
        self.dropout3 = nn.Dropout(dropout)
        self.norm2 = nn.LayerNorm(d_model)

    @staticmethod
    def with_pos_embed(tensor, pos):
        return tensor if pos is None else tensor + pos

    def forward_ffn(self, src):
        return self._ffn(_Stream.start(src, None), None, False).x

    def _ffn(self, st, pos, want_pos):
        h = self.dropout2(_act_linear(st.x_lp, self.linear1, self.activation))
        return _norm_step(st.x, self.dropout3(linear(h, self.linear2)), self.norm2, pos, want_pos)

    def _forward_stream(self, st, pos, reference_points, spatial_shapes, padding_mask, last):
        # self attention: query = src + pos, value source = src   (reference :293-296)
        src2 = self.self_attn(st.xpos_lp, reference_points, st.x_lp, spatial_shapes, padding_mask)
        st = _norm_step(st.x, self.dropout1(src2), self.norm1, None, False)
        # ffn (:282-287); the closing norm also prepares src + pos for the next layer's query
        return self._ffn(st, pos, pos is not None and not last)

    def forward(self, src, pos, reference_points, spatial_shapes, padding_mask=None):
        return self._forward_stream(_Stream.start(src, pos), pos, reference_points, spatial_shapes,
                                    padding_mask, True).x


class DeformableTransformerEncoder(nn.Module):
    def __init__(self, encoder_layer, num_layers):
        super().__init__()
        self.layers = _get_clones(encoder_layer, num_layers)
        self.num_layers = num_layers

    @staticmethod
    def get_reference_points(spatial_shapes, valid_ratios, device):
        """Pixel centres of every level, normalised by the valid extent, for every level (reference :308-321)."""
        per_level = []
        for lvl, (gy, gx, H_, W_) in enumerate(_level_grids(spatial_shapes, device)):
            ref_y = gy[None] / (valid_ratios[:, None, lvl, 1] * H_)
            ref_x = gx[None] / (valid_ratios[:, None, lvl, 0] * W_)
            per_level.append(torch.stack((ref_x, ref_y), -1))
        reference_points = torch.cat(per_level, 1)
        return reference_points[:, :, None] * valid_ratios[:, None]

    def forward(self, src, spatial_shapes, valid_ratios, pos=None, padding_mask=None, return_stream=False):
        reference_points = self.get_reference_points(spatial_shapes, valid_ratios, device=src.device)
        st = _Stream.start(src, pos)
        for i, layer in enumerate(self.layers):
            st = layer._forward_stream(st, pos, reference_points, spatial_shapes, padding_mask,
                                       i == self.num_layers - 1)
        return st if return_stream else st.x


class DeformableTransformerDecoderLayer(nn.Module):
    def __init__(self, d_model=256, d_ffn=1024, dropout=0.1, activation="relu",
                 n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        # cross attention
        self.cross_attn = MSDeformAttn(d_model, n_levels, n_heads, n_points)
        self.dropout1 = nn.Dropout(dropout)
        self.norm1 = nn.LayerNorm(d_model)
        # self attention
        self.self_attn = nn.MultiheadAttention(d_model, n_heads, dropout=dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.norm2 = nn.LayerNorm(d_model)
        # ffn
        self.linear1 = nn.Linear(d_model, d_ffn)
        self.activation = _get_activation_fn(activation)
        self.dropout3 = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_ffn, d_model)
        self.dropout4 = nn.Dropout(dropout)
        self.norm3 = nn.LayerNorm(d_model)

    @staticmethod
    def with_pos_embed(tensor, pos):
        return tensor if pos is None else tensor + pos

    def forward_ffn(self, tgt):
        return self._ffn(_Stream.start(tgt, None), None, False).x

    def _ffn(self, st, pos, want_pos):
        h = self.dropout3(_act_linear(st.x_lp, self.linear1, self.activation))
        return _norm_step(st.x, self.dropout4(linear(h, self.linear2)), self.norm3, pos, want_pos)

    def _forward_stream(self, st, query_pos, reference_points, src, src_spatial_shapes,
                        src_padding_mask, query_attn_mask, last):
        # self attention over the queries: q = k = tgt + query_pos, v = tgt   (reference :367-376)
        mha = self.self_attn
        if mha._qkv_same_embed_dim and mha.in_proj_bias is not None and mha.bias_k is None and not mha.add_zero_attn:
            tgt2 = _self_attention(mha, st.xpos_lp, st.x_lp, query_attn_mask)
        else:
            q = st.xpos_lp.transpose(0, 1)
            tgt2 = mha(q, q, st.x_lp.transpose(0, 1), key_padding_mask=query_attn_mask,
                       need_weights=False)[0].transpose(0, 1)
        st = _norm_step(st.x, self.dropout2(tgt2), self.norm2, query_pos, query_pos is not None)
        # cross attention into the encoder memory (:378-383)
        tgt2 = self.cross_attn(st.xpos_lp, reference_points, src, src_spatial_shapes,
                               src_padding_mask, query_attn_mask)
        st = _norm_step(st.x, self.dropout1(tgt2), self.norm1, None, False)
        # ffn (:361-365); the closing norm prepares tgt + query_pos for the next layer
        return self._ffn(st, query_pos, query_pos is not None and not last)

    def forward(self, tgt, query_pos, reference_points, src, src_spatial_shapes, src_padding_mask=None,
                query_attn_mask=None):
        return self._forward_stream(_Stream.start(tgt, query_pos), query_pos, reference_points, src,
                                    src_spatial_shapes, src_padding_mask, query_attn_mask, True).x


class DeformableTransformerDecoder(nn.Module):
    def __init__(self, decoder_layer, num_layers, return_intermediate=False):
        super().__init__()
        self.layers = _get_clones(decoder_layer, num_layers)
        self.num_layers = num_layers
        self.return_intermediate = return_intermediate
        # set by the detection head for iterative box refinement / two-stage (reference :396-398)
        self.bbox_embed = None
        self.class_embed = None

    def forward(self, tgt, reference_points, src, src_spatial_shapes, src_valid_ratios,
                query_pos=None, src_padding_mask=None, query_attn_mask=None):
        """src may be the encoder's output tensor or the _Stream returned by
        DeformableTransformerEncoder.forward(..., return_stream=True) (its low-precision copy is reused)."""
        memory = src.x_lp if isinstance(src, _Stream) else src
        st = _Stream.start(tgt, query_pos)
        intermediate, intermediate_reference_points = [], []
        for lid, layer in enumerate(self.layers):
            if reference_points.shape[-1] == 4:
                ratios = torch.cat([src_valid_ratios, src_valid_ratios], -1)
            else:
                assert reference_points.shape[-1] == 2
                ratios = src_valid_ratios
            reference_points_input = reference_points[:, :, None] * ratios[:, None]
            refine = self.bbox_embed is not None
            # a refined layer's output feeds bbox_embed and the next layer alike: keep tgt + query_pos coming
            st = layer._forward_stream(st, query_pos, reference_points_input, memory, src_spatial_shapes,
                                       src_padding_mask, query_attn_mask, lid == self.num_layers - 1)
            output = st.x

            if refine:      # iterative bounding box refinement (reference :413-424)
                tmp = self.bbox_embed[lid](output)
                if reference_points.shape[-1] == 4:
                    new_reference_points = (tmp + inverse_sigmoid(reference_points)).sigmoid()
                else:
                    new_reference_points = tmp
                    new_reference_points[..., :2] = tmp[..., :2] + inverse_sigmoid(reference_points)
                    new_reference_points = new_reference_points.sigmoid()
                reference_points = new_reference_points.detach()

            if self.return_intermediate:
                intermediate.append(output)
                intermediate_reference_points.append(reference_points)

        if self.return_intermediate:
            return torch.stack(intermediate), torch.stack(intermediate_reference_points)
        return output, reference_points
