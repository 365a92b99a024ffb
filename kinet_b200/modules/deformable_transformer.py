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
import math

import weakref

import torch
import torch.nn.functional as F
from torch import nn

from .._cache import PerTensorCache
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


_grids = PerTensorCache()   # spatial_shapes tensor object -> per-level pixel-centre grids, keyed additionally by device


def _level_grids(spatial_shapes, device):
    """Flattened pixel-centre coordinates (y, x) of every level.  Reading H and W out of a CUDA
    ``spatial_shapes`` tensor is a device->host sync (the reference pays it on every forward, :311-314, and it
    is illegal while a CUDA graph is being captured), so the grids are remembered for the tensor objects
    (identity + version counter) they were built from; an eager warm-up run with the same tensor fills them."""
    is_t = torch.is_tensor(spatial_shapes)
    if is_t:
        hit = _grids.get(spatial_shapes, str(device))
        if hit is not None:
            return hit
    out = []
    for H_, W_ in (spatial_shapes.tolist() if is_t else spatial_shapes):
        H_, W_ = int(H_), int(W_)
        ys = torch.linspace(0.5, H_ - 0.5, H_, dtype=torch.float32, device=device)
        xs = torch.linspace(0.5, W_ - 0.5, W_, dtype=torch.float32, device=device)
        gy, gx = torch.meshgrid(ys, xs, indexing="ij")
        out.append((gy.reshape(-1), gx.reshape(-1), H_, W_))
    if is_t:
        _grids.put(spatial_shapes, out, str(device))
    return out


def _act_linear(x, layer, activation, drop=None):
    """drop(activation(layer(x))); ReLU rides in the GEMM's epilogue (and, in the backward, in the bias-gradient pass,
    together with the dropout's backward: no mask tensor, functions/layer_ops.py::dropout_)."""
    p = drop.p if (drop is not None and drop.training) else 0.0
    if activation is F.relu:
        return linear(x, layer, relu=True, drop_p=p)
    y = activation(linear(x, layer))
    return drop(y) if drop is not None else y


class _skip_output_bias_grad:
    """While active, ``attn`` (an MSDeformAttn) leaves the bias gradient of its output_proj to the add + LayerNorm
    step that consumes its output (``_norm_step(..., branch_bias=attn.output_proj.bias)``)."""

    def __init__(self, attn):
        self.attn = attn

    def __enter__(self):
        self.attn._kinet_skip_output_bias_grad = True

    def __exit__(self, *exc):
        self.attn._kinet_skip_output_bias_grad = False


def _self_attention(mha, q_in, v_in, key_padding_mask, skip_out_bias_grad=False):
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
    return linear(o.transpose(1, 2).reshape(N, Lq, C), mha.out_proj, skip_bias_grad=skip_out_bias_grad)


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


def _norm_step(x, branch, norm, pos, want_pos, drop=None, branch_bias=None):
    """x <- norm(x + drop(branch)); returns the next _Stream (copies only where the stream dtype differs).
    drop: the layer's nn.Dropout for this branch -- applied INSIDE the fused kernel (no dropout kernel, no mask
    tensor; the backward regenerates the mask).  branch_bias: bias of the Linear layer that produced ``branch`` when
    that layer was told to skip its bias gradient (it then comes out of this step's backward)."""
    lp = _lp_dtype(x)
    same = lp == x.dtype
    p = drop.p if (drop is not None and drop.training) else 0.0
    if branch_bias is not None and not (torch.is_grad_enabled() and branch_bias.requires_grad):
        branch_bias = None
    y, y_lp, ypos = add_layer_norm(x, branch, norm, pos=pos if want_pos else None, lp_dtype=lp, want_lp=not same,
                                   drop_p=p, branch_bias=branch_bias)
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
        h = _act_linear(st.x_lp, self.linear1, self.activation, self.dropout2)
        # dropout3 and linear2's bias gradient ride in the fused add + LayerNorm step
        return _norm_step(st.x, linear(h, self.linear2, skip_bias_grad=True), self.norm2, pos, want_pos,
                          drop=self.dropout3, branch_bias=self.linear2.bias)

    def _forward_stream(self, st, pos, reference_points, spatial_shapes, padding_mask, last):
        # self attention: query = src + pos, value source = src   (reference :293-296)
        with _skip_output_bias_grad(self.self_attn):
            src2 = self.self_attn(st.xpos_lp, reference_points, st.x_lp, spatial_shapes, padding_mask)
        st = _norm_step(st.x, src2, self.norm1, None, False, drop=self.dropout1,
                        branch_bias=self.self_attn.output_proj.bias)
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
        h = _act_linear(st.x_lp, self.linear1, self.activation, self.dropout3)
        return _norm_step(st.x, linear(h, self.linear2, skip_bias_grad=True), self.norm3, pos, want_pos,
                          drop=self.dropout4, branch_bias=self.linear2.bias)

    def _forward_stream(self, st, query_pos, reference_points, src, src_spatial_shapes,
                        src_padding_mask, query_attn_mask, last):
        # self attention over the queries: q = k = tgt + query_pos, v = tgt   (reference :367-376)
        mha = self.self_attn
        out_bias = None
        if mha._qkv_same_embed_dim and mha.in_proj_bias is not None and mha.bias_k is None and not mha.add_zero_attn:
            tgt2 = _self_attention(mha, st.xpos_lp, st.x_lp, query_attn_mask, skip_out_bias_grad=True)
            out_bias = mha.out_proj.bias
        else:
            q = st.xpos_lp.transpose(0, 1)
            tgt2 = mha(q, q, st.x_lp.transpose(0, 1), key_padding_mask=query_attn_mask,
                       need_weights=False)[0].transpose(0, 1)
        st = _norm_step(st.x, tgt2, self.norm2, query_pos, query_pos is not None, drop=self.dropout2,
                        branch_bias=out_bias)
        # cross attention into the encoder memory (:378-383)
        with _skip_output_bias_grad(self.cross_attn):
            tgt2 = self.cross_attn(st.xpos_lp, reference_points, src, src_spatial_shapes,
                                   src_padding_mask, query_attn_mask)
        st = _norm_step(st.x, tgt2, self.norm1, None, False, drop=self.dropout1,
                        branch_bias=self.cross_attn.output_proj.bias)
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


_mask_lists = {}        # (id, version) of every per-level mask tensor -> True when none of them masks anything


def _remember_no_padding(masks, mask_flatten):
    """The flattened padding mask is a new tensor on every forward, so MSDeformAttn's per-tensor "masks nothing"
    answer (one device->host read) would be asked again every frame -- and cannot be asked at all while a CUDA graph
    is captured.  Callers that pass the SAME per-level mask tensors every time (a tracker on a fixed-size stream, a
    graphed training step with static input buffers) get the answer from the first eager forward."""
    from .ms_deform_attn import _mask_empty, _no_padding
    if any(m.is_inference() for m in masks):
        return
    key = tuple((id(m), m._version) for m in masks)
    hit = _mask_lists.get(key)
    if hit is None:
        if mask_flatten.is_cuda and torch.cuda.is_current_stream_capturing():
            return
        if len(_mask_lists) > 16:
            _mask_lists.clear()
        hit = _mask_lists[key] = (_no_padding(mask_flatten), [weakref.ref(m) for m in masks])
    elif any(r() is not m for r, m in zip(hit[1], masks)):       # ids were reused by other tensors
        _mask_lists.pop(key, None)
        return
    _mask_empty.put(mask_flatten, hit[0])


_shape_tensors = {}     # (tuple of (H, W), device) -> int64 tensor, so every forward hands the SAME tensor object down


def _shapes_tensor(shapes, device):
    """The (L, 2) int64 level-shape tensor for a pyramid.  The reference builds a fresh tensor on every forward
    (:155), which defeats the per-tensor caches above and in MSDeformAttn (shape check, level grids, offset
    divisors) and costs a host->device copy per frame; pyramids repeat, so the tensor is kept per shape list."""
    key = (tuple(shapes), str(device))
    t = _shape_tensors.get(key)
    if t is None:
        if len(_shape_tensors) > 64:
            _shape_tensors.clear()
        t = torch.as_tensor(shapes, dtype=torch.long, device=device)
        _shape_tensors[key] = t
    return t


class DeformableTransformer(nn.Module):
    """The reference's DeformableTransformer (:19-257): flattens the feature pyramid, runs the encoder (once, or
    once per frame with ``multi_frame_attention_separate_encoder``), builds the decoder's queries (learned object
    queries, optionally preceded by the previous frame's track queries from ``targets``; or two-stage proposals)
    and runs the decoder.  Same constructor arguments, parameter names (``level_embed``, ``reference_points``,
    ``enc_output*``, ``pos_trans*``) and return tuple."""

    def __init__(self, d_model=256, nhead=8, num_encoder_layers=6, num_decoder_layers=6, dim_feedforward=1024,
                 dropout=0.1, activation="relu", return_intermediate_dec=False, num_feature_levels=4,
                 dec_n_points=4, enc_n_points=4, two_stage=False, two_stage_num_proposals=300,
                 multi_frame_attention_separate_encoder=False):
        super().__init__()
        self.d_model = d_model
        self.nhead = nhead
        self.two_stage = two_stage
        self.two_stage_num_proposals = two_stage_num_proposals
        self.num_feature_levels = num_feature_levels
        self.multi_frame_attention_separate_encoder = multi_frame_attention_separate_encoder

        enc_levels = num_feature_levels // 2 if multi_frame_attention_separate_encoder else num_feature_levels
        self.encoder = DeformableTransformerEncoder(
            DeformableTransformerEncoderLayer(d_model, dim_feedforward, dropout, activation, enc_levels, nhead,
                                              enc_n_points), num_encoder_layers)
        self.decoder = DeformableTransformerDecoder(
            DeformableTransformerDecoderLayer(d_model, dim_feedforward, dropout, activation, num_feature_levels, nhead,
                                              dec_n_points), num_decoder_layers, return_intermediate_dec)
        self.level_embed = nn.Parameter(torch.Tensor(num_feature_levels, d_model))
        if two_stage:
            self.enc_output = nn.Linear(d_model, d_model)
            self.enc_output_norm = nn.LayerNorm(d_model)
            self.pos_trans = nn.Linear(d_model * 2, d_model * 2)
            self.pos_trans_norm = nn.LayerNorm(d_model * 2)
        else:
            self.reference_points = nn.Linear(d_model, 2)
        self._reset_parameters()

    def _reset_parameters(self):
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)
        for m in self.modules():
            if isinstance(m, MSDeformAttn):
                m._reset_parameters()
        if not self.two_stage:
            nn.init.xavier_uniform_(self.reference_points.weight.data, gain=1.0)
            nn.init.constant_(self.reference_points.bias.data, 0.)
        nn.init.normal_(self.level_embed)

    @staticmethod
    def get_valid_ratio(mask):
        """Fraction of each image that is not padding, (w, h) order, from a (N, H, W) mask (True = padding)."""
        _, H, W = mask.shape
        valid_h = (~mask[:, :, 0]).sum(1).float() / H
        valid_w = (~mask[:, 0, :]).sum(1).float() / W
        return torch.stack([valid_w, valid_h], -1)

    def get_proposal_pos_embed(self, proposals):
        """Sine embedding of two-stage proposals (N, L, 4) -> (N, L, 512) (reference :70-83)."""
        num_pos_feats, temperature = 128, 10000
        dim_t = torch.arange(num_pos_feats, dtype=torch.float32, device=proposals.device)
        dim_t = temperature ** (2 * (dim_t // 2) / num_pos_feats)
        pos = (proposals.sigmoid() * (2 * math.pi))[:, :, :, None] / dim_t
        return torch.stack((pos[:, :, :, 0::2].sin(), pos[:, :, :, 1::2].cos()), dim=4).flatten(2)

    def gen_encoder_output_proposals(self, memory, memory_padding_mask, spatial_shapes):
        """Two-stage: one box proposal per encoder token (pixel centre, level-dependent size) in logit space, invalid
        ones pushed to +inf, and the projected memory they are scored from (reference :85-117).  ``spatial_shapes``
        is the Python list of (H, W)."""
        N_ = memory.shape[0]
        proposals, cur = [], 0
        for lvl, (H_, W_) in enumerate(spatial_shapes):
            m = memory_padding_mask[:, cur:cur + H_ * W_].view(N_, H_, W_, 1)
            valid_h = (~m[:, :, 0, 0]).sum(1)
            valid_w = (~m[:, 0, :, 0]).sum(1)
            gy, gx = torch.meshgrid(torch.linspace(0, H_ - 1, H_, dtype=torch.float32, device=memory.device),
                                    torch.linspace(0, W_ - 1, W_, dtype=torch.float32, device=memory.device), indexing="ij")
            grid = torch.stack([gx, gy], -1)
            scale = torch.stack([valid_w, valid_h], 1).view(N_, 1, 1, 2)
            grid = (grid[None].expand(N_, -1, -1, -1) + 0.5) / scale
            wh = torch.ones_like(grid) * 0.05 * (2.0 ** lvl)
            proposals.append(torch.cat((grid, wh), -1).view(N_, -1, 4))
            cur += H_ * W_
        out = torch.cat(proposals, 1)
        valid = ((out > 0.01) & (out < 0.99)).all(-1, keepdim=True)
        out = torch.log(out / (1 - out))
        out = out.masked_fill(memory_padding_mask.unsqueeze(-1), float("inf")).masked_fill(~valid, float("inf"))
        mem = memory.masked_fill(memory_padding_mask.unsqueeze(-1), 0.0).masked_fill(~valid, 0.0)
        return self.enc_output_norm(self.enc_output(mem)), out

    def forward(self, srcs, masks, pos_embeds, query_embed=None, targets=None):
        """srcs / pos_embeds: per level (N, C, H, W); masks: per level (N, H, W) bool, True = padding;
        query_embed (Lq, 2C) = [query_pos | tgt] (required unless two_stage); targets: optional list of dicts with
        'track_query_hs_embeds' (T, C) and 'track_query_boxes' (T, 4) -- the previous frame's tracks, which become
        the first T decoder queries.  -> (hs, memory, init_reference, inter_references, enc_class, enc_coord)."""
        assert self.two_stage or query_embed is not None
        shapes, src_l, mask_l, pos_l = [], [], [], []
        for lvl, (src, mask, pos) in enumerate(zip(srcs, masks, pos_embeds)):
            shapes.append((int(src.shape[2]), int(src.shape[3])))
            src_l.append(src.flatten(2).transpose(1, 2))
            mask_l.append(mask.flatten(1))
            pos_l.append(pos.flatten(2).transpose(1, 2) + self.level_embed[lvl].view(1, 1, -1))
        src_flatten, mask_flatten, pos_flatten = torch.cat(src_l, 1), torch.cat(mask_l, 1), torch.cat(pos_l, 1)
        _remember_no_padding(masks, mask_flatten)
        device = src_flatten.device
        spatial_shapes = _shapes_tensor(shapes, device)
        valid_ratios = torch.stack([self.get_valid_ratio(m) for m in masks], 1)

        if self.multi_frame_attention_separate_encoder:
            # levels [0, L/2) are the current frame's, [L/2, L) the previous frame's; one encoder, run per frame
            half, hl = src_flatten.shape[1] // 2, self.num_feature_levels // 2
            prev_memory = self.encoder(src_flatten[:, :half], _shapes_tensor(shapes[:hl], device), valid_ratios[:, :hl],
                                       pos_flatten[:, :half], mask_flatten[:, :half])
            memory = self.encoder(src_flatten[:, half:], _shapes_tensor(shapes[hl:], device), valid_ratios[:, hl:],
                                  pos_flatten[:, half:], mask_flatten[:, half:])
            memory = torch.cat([memory, prev_memory], 1)
            memory_for_decoder = memory
        else:
            stream = self.encoder(src_flatten, spatial_shapes, valid_ratios, pos_flatten, mask_flatten,
                                  return_stream=True)
            memory, memory_for_decoder = stream.x, stream        # the decoder reuses the low-precision copy

        bs, _, c = memory.shape
        query_attn_mask = None
        enc_outputs_class = enc_outputs_coord_unact = None
        if self.two_stage:
            output_memory, output_proposals = self.gen_encoder_output_proposals(memory, mask_flatten, shapes)
            enc_outputs_class = self.decoder.class_embed[self.decoder.num_layers](output_memory)
            enc_outputs_coord_unact = self.decoder.bbox_embed[self.decoder.num_layers](output_memory) + output_proposals
            top = torch.topk(enc_outputs_class[..., 0], self.two_stage_num_proposals, dim=1)[1]
            top_coords = torch.gather(enc_outputs_coord_unact, 1, top.unsqueeze(-1).repeat(1, 1, 4)).detach()
            reference_points = top_coords.sigmoid()
            query_embed, tgt = torch.split(self.pos_trans_norm(self.pos_trans(self.get_proposal_pos_embed(top_coords))),
                                           c, dim=2)
        else:
            query_embed, tgt = torch.split(query_embed, c, dim=1)
            query_embed = query_embed.unsqueeze(0).expand(bs, -1, -1)
            tgt = tgt.unsqueeze(0).expand(bs, -1, -1)
            reference_points = self.reference_points(query_embed).sigmoid()
            if targets is not None and "track_query_hs_embeds" in targets[0]:
                # track queries: content = the track's last decoder output, no positional part, reference = box centre
                prev_hs = torch.stack([t["track_query_hs_embeds"] for t in targets])
                prev_boxes = torch.stack([t["track_query_boxes"] for t in targets])
                query_embed = torch.cat([torch.zeros_like(prev_hs), query_embed], dim=1)
                tgt = torch.cat([prev_hs, tgt], dim=1)
                reference_points = torch.cat([prev_boxes[..., :2], reference_points], dim=1)
        init_reference_out = reference_points

        hs, inter_references = self.decoder(tgt, reference_points, memory_for_decoder, spatial_shapes, valid_ratios,
                                            query_embed, mask_flatten, query_attn_mask)
        return hs, memory, init_reference_out, inter_references, enc_outputs_class, enc_outputs_coord_unact
