"""The literal KineT (`kine: True`) transformer -- SURVEY.md section 8 row f4 -- mirroring the reference's
src/trackformer/models/transformer.py:

    KinematicTransformer       :85-142       DualKinematicTransformer   :145-185
    DualKinematicEncoder       :188-242      TransformerEncoder         :245-266
    TransformerDecoder         :269-326      TransformerEncoderLayer    :329-385
    TransformerDecoderLayer    :388-469      IntertwinedBranch          :472-488

Same constructor arguments, sub-module / parameter names (``self_attn``, ``multihead_attn``, ``linear1/2``,
``norm1..3``, ``transformer_det`` ...; checkpoints load unchanged), forward signatures, tensor layouts of the inputs
and outputs (the reference's layers are sequence-first ``(T, BS, C)``) and arithmetic.  It contains no MSDeformAttn:
the tokens are detection boxes / tracklets, <= 150 + T of them, model width 288 (cfgs/train_kinet.yaml:17-27).  What
changes is how it runs on the GPU:

  * the attention core of every ``nn.MultiheadAttention`` is ONE hand-written kernel (csrc/small_mha.cu: fp32
    arithmetic, key-padding mask, online softmax; backward = two kernels without atomics) instead of the library's
    chain of small launches -- heads of 36 channels are what library attention kernels handle worst;
  * every ``x = norm(x + dropout(branch))`` and the ``with_pos_embed`` behind it is the fused add + LayerNorm kernel
    of the deformable layers (functions/layer_ops.py), which also emits the low-precision copies the next GEMMs read
    under CUDA autocast; ReLU rides in the GEMM epilogue; bias gradients are fused column sums;
  * internally everything is batch-first and contiguous; the sequence-first views the reference returns are
    ``transpose`` views of those tensors.

The projections stay cuBLAS GEMMs.  Not on the fused path, computed with plain PyTorch CUDA ops instead:
``normalize_before=True`` (pre-norm; the reference's configurations use post-norm, cfgs/train.yaml:62) and
attention-weight dropout in training mode (``dropout > 0`` and ``.train()``: F.scaled_dot_product_attention with
``dropout_p``).  ``attn_mask`` arguments (``src_mask`` / ``tgt_mask`` / ``memory_mask``) must be None -- the reference
never passes one on this path.  CPU tensors raise from the kernels' wrappers.
"""
import torch
import torch.nn.functional as F
from torch import nn

from ..functions.layer_ops import add_layer_norm, inference_mirror, linear, linear_params
from ..functions.small_mha import small_mha, supported as _mha_supported
from .deformable_transformer import _Stream, _act_linear, _get_activation_fn, _get_clones, _lp_dtype, _norm_step


def _bf(t):
    """sequence-first (T, BS, C) -> batch-first contiguous (BS, T, C)."""
    return None if t is None else t.transpose(0, 1).contiguous()


def _attention(mha, q_in, k_in, v_in, key_padding_mask, self_attention):
    """nn.MultiheadAttention(q, k, v, key_padding_mask=...)[0] on batch-first inputs, from the module's own
    parameters: the input projections through ``linear_params`` (q and k in one GEMM when they share their input),
    the attention core through ``small_mha``, ``out_proj``."""
    if not (mha._qkv_same_embed_dim and mha.in_proj_bias is not None and mha.bias_k is None and not mha.add_zero_attn):
        raise RuntimeError("kinet_b200: nn.MultiheadAttention variant not supported (kdim / vdim / bias_k / add_zero_attn)")
    N, Lq, C = q_in.shape
    Lk = k_in.shape[1]
    H = mha.num_heads
    D = C // H
    w, b = mha.in_proj_weight, mha.in_proj_bias
    mirror = inference_mirror(mha, w, b, q_in)

    def part(lo, hi):
        return None if mirror is None else (mirror[0][lo:hi], mirror[1][lo:hi])
    if self_attention:
        qk = linear_params(q_in, w[:2 * C], b[:2 * C], False, part(0, 2 * C))
        q, k = qk[..., :C], qk[..., C:]
    else:
        q = linear_params(q_in, w[:C], b[:C], False, part(0, C))
        k = linear_params(k_in, w[C:2 * C], b[C:2 * C], False, part(C, 2 * C))
    v = linear_params(v_in, w[2 * C:], b[2 * C:], False, part(2 * C, 3 * C))
    q4, k4, v4 = q.unflatten(-1, (H, D)), k.unflatten(-1, (H, D)), v.unflatten(-1, (H, D))
    drop = mha.dropout if mha.training else 0.0
    if drop > 0.0 or not _mha_supported(q4, k4, v4):
        # attention-weight dropout (training) or heads wider than the kernel's limit: the library kernel
        mask = None if key_padding_mask is None else ~key_padding_mask[:, None, None, :]
        o = F.scaled_dot_product_attention(q4.transpose(1, 2), k4.transpose(1, 2), v4.transpose(1, 2),
                                           attn_mask=mask, dropout_p=drop).transpose(1, 2).reshape(N, Lq, C)
    else:
        o = small_mha(q4, k4, v4, key_padding_mask)
    return linear(o, mha.out_proj)


def _two_branches(fn_a, fn_b, device):
    """(fn_a(), fn_b()) with fn_b on a second CUDA stream.  The detection and the metadata transformer do not see each
    other until the IntertwinedBranch blocks, and every kernel of either occupies a fraction of the GPU (<= 150 CTAs,
    2-5 us): side by side -- as two parallel branches of a captured CUDA graph, or two streams in eager mode -- they take
    little more than one of them.  Inference only (no autograd graph crosses the streams here)."""
    if device.type != "cuda" or torch.is_grad_enabled():
        return fn_a(), fn_b()
    main = torch.cuda.current_stream(device)
    side = _side_streams.get(device.index)
    if side is None:
        side = _side_streams[device.index] = torch.cuda.Stream(device=device)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        b = fn_b()
    a = fn_a()
    main.wait_stream(side)
    for t in (b if isinstance(b, (tuple, list)) else (b,)):
        if torch.is_tensor(t):
            t.record_stream(main)           # allocated on the side stream, consumed (and freed) on the main one
    return a, b


_side_streams = {}


def _check_no_attn_mask(*masks):
    if any(m is not None for m in masks):
        raise RuntimeError("kinet_b200: attn_mask is not supported on the kine path (the reference passes None)")


class TransformerEncoderLayer(nn.Module):
    def __init__(self, d_model, nhead, dim_feedforward=2048, dropout=0.1, activation="relu", normalize_before=False):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d_model, nhead, dropout=dropout)
        self.linear1 = nn.Linear(d_model, dim_feedforward)
        self.dropout = nn.Dropout(dropout)
        self.linear2 = nn.Linear(dim_feedforward, d_model)
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.activation = _get_activation_fn(activation)
        self.normalize_before = normalize_before

    @staticmethod
    def with_pos_embed(tensor, pos):
        return tensor if pos is None else tensor + pos

    # ---- batch-first streams (what the containers call) ----
    def _post_stream(self, st, pos, mask, last):
        src2 = _attention(self.self_attn, st.xpos_lp, st.xpos_lp, st.x_lp, mask, True)          # :331-333
        # (the branch dropouts are generated inside the fused add + LayerNorm kernels, the FFN's inner one without a mask)
        st = _norm_step(st.x, src2, self.norm1, None, False, drop=self.dropout1)                 # :334-335
        h = _act_linear(st.x_lp, self.linear1, self.activation, self.dropout)                    # :336
        return _norm_step(st.x, linear(h, self.linear2, skip_bias_grad=True), self.norm2, pos,   # :337-338
                          pos is not None and not last, drop=self.dropout2, branch_bias=self.linear2.bias)

    def _pre_bf(self, src, pos, mask):
        src2 = self.norm1(src)                                                                   # :345-352
        q = self.with_pos_embed(src2, pos)
        src = src + self.dropout1(_attention(self.self_attn, q, q, src2, mask, True))
        src2 = self.norm2(src)
        return src + self.dropout2(self.linear2(self.dropout(self.activation(self.linear1(src2)))))

    def _forward_bf(self, st, pos, mask, last):
        if self.normalize_before:
            return _Stream.start(self._pre_bf(st.x, pos, mask), None if last else pos)
        return self._post_stream(st, pos, mask, last)

    # ---- the reference's sequence-first interface ----
    def forward(self, src, src_mask=None, src_key_padding_mask=None, pos=None):
        _check_no_attn_mask(src_mask)
        posb = _bf(pos)
        return self._forward_bf(_Stream.start(_bf(src), posb), posb, src_key_padding_mask, True).x.transpose(0, 1)


class TransformerDecoderLayer(nn.Module):
    def __init__(self, d_model, nhead, dim_feedforward=2048, dropout=0.1, activation="relu", normalize_before=False):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d_model, nhead, dropout=dropout)
        self.multihead_attn = nn.MultiheadAttention(d_model, nhead, dropout=dropout)
        self.linear1 = nn.Linear(d_model, dim_feedforward)
        self.dropout = nn.Dropout(dropout)
        self.linear2 = nn.Linear(dim_feedforward, d_model)
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self.norm3 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.dropout3 = nn.Dropout(dropout)
        self.activation = _get_activation_fn(activation)
        self.normalize_before = normalize_before

    @staticmethod
    def with_pos_embed(tensor, pos):
        return tensor if pos is None else tensor + pos

    def _post_stream(self, st, mem_lp, mempos_lp, query_pos, tgt_mask_kp, mem_mask_kp, last):
        tgt2 = _attention(self.self_attn, st.xpos_lp, st.xpos_lp, st.x_lp, tgt_mask_kp, True)    # :393-395
        st = _norm_step(st.x, tgt2, self.norm1, query_pos, query_pos is not None, drop=self.dropout1)  # :396-397
        tgt2 = _attention(self.multihead_attn, st.xpos_lp, mempos_lp, mem_lp, mem_mask_kp, False)  # :398-401
        st = _norm_step(st.x, tgt2, self.norm2, None, False, drop=self.dropout2)                  # :402-403
        h = _act_linear(st.x_lp, self.linear1, self.activation, self.dropout)                     # :404
        return _norm_step(st.x, linear(h, self.linear2, skip_bias_grad=True), self.norm3, query_pos,   # :405-406
                          query_pos is not None and not last, drop=self.dropout3, branch_bias=self.linear2.bias)

    def _pre_bf(self, tgt, memory, mempos, query_pos, tgt_mask_kp, mem_mask_kp):
        tgt2 = self.norm1(tgt)                                                                    # :416-430
        q = self.with_pos_embed(tgt2, query_pos)
        tgt = tgt + self.dropout1(_attention(self.self_attn, q, q, tgt2, tgt_mask_kp, True))
        tgt2 = self.norm2(tgt)
        tgt = tgt + self.dropout2(_attention(self.multihead_attn, self.with_pos_embed(tgt2, query_pos), mempos, memory,
                                             mem_mask_kp, False))
        tgt2 = self.norm3(tgt)
        return tgt + self.dropout3(self.linear2(self.dropout(self.activation(self.linear1(tgt2)))))

    def _forward_bf(self, st, memory, mem_lp, mempos, mempos_lp, query_pos, tgt_mask_kp, mem_mask_kp, last):
        if self.normalize_before:
            return _Stream.start(self._pre_bf(st.x, memory, mempos, query_pos, tgt_mask_kp, mem_mask_kp),
                                 None if last else query_pos)
        return self._post_stream(st, mem_lp, mempos_lp, query_pos, tgt_mask_kp, mem_mask_kp, last)

    def forward(self, tgt, memory, tgt_mask=None, memory_mask=None, tgt_key_padding_mask=None,
                memory_key_padding_mask=None, pos=None, query_pos=None):
        _check_no_attn_mask(tgt_mask, memory_mask)
        memb, posb, qposb = _bf(memory), _bf(pos), _bf(query_pos)
        mem = _Stream.start(memb, posb)
        return self._forward_bf(_Stream.start(_bf(tgt), qposb), memb, mem.x_lp, self.with_pos_embed(memb, posb),
                                mem.xpos_lp, qposb, tgt_key_padding_mask, memory_key_padding_mask,
                                True).x.transpose(0, 1)


class TransformerEncoder(nn.Module):
    def __init__(self, encoder_layer, num_layers, norm=None):
        super().__init__()
        self.layers = _get_clones(encoder_layer, num_layers)
        self.num_layers = num_layers
        self.norm = norm

    def _forward_bf(self, src, mask, pos):
        st = _Stream.start(src, pos)
        for i, layer in enumerate(self.layers):
            st = layer._forward_bf(st, pos, mask, i == self.num_layers - 1)
        return self.norm(st.x) if self.norm is not None else st.x

    def forward(self, src, mask=None, src_key_padding_mask=None, pos=None):
        _check_no_attn_mask(mask)
        return self._forward_bf(_bf(src), src_key_padding_mask, _bf(pos)).transpose(0, 1)


class TransformerDecoder(nn.Module):
    def __init__(self, decoder_layer, encoder_layer, num_layers, norm=None, return_intermediate=False,
                 track_attention=False):
        super().__init__()
        self.layers = _get_clones(decoder_layer, num_layers)
        self.num_layers = num_layers
        self.norm = norm
        self.return_intermediate = return_intermediate
        self.track_attention = track_attention
        if self.track_attention:
            self.layers_track_attention = _get_clones(encoder_layer, num_layers)

    def _forward_bf(self, tgt, memory, tgt_mask_kp, mem_mask_kp, pos, query_pos):
        """Batch-first.  Returns (norm(output), output): (layers, BS, Lq, C) each with return_intermediate."""
        mem = _Stream.start(memory, pos)
        mempos = None                                              # fp32 memory + pos: the pre-norm layers' key input
        if self.layers[0].normalize_before:
            mempos = memory if pos is None else memory + pos
        track_query_pos = None
        if self.track_attention:                                   # :290-292 (the last 100 queries are object queries)
            track_query_pos = query_pos[:, :-100].clone()
            query_pos = query_pos.clone()
            query_pos[:, :-100] = 0.0
        st = _Stream.start(tgt, query_pos)
        intermediate = []
        for i, layer in enumerate(self.layers):
            last = i == self.num_layers - 1
            if self.track_attention:                               # :295-305
                track = self.layers_track_attention[i]._forward_bf(
                    _Stream.start(st.x[:, :-100].contiguous(), track_query_pos), track_query_pos, tgt_mask_kp, True).x
                st = _Stream.start(torch.cat([track, st.x[:, -100:]], 1), query_pos)
            st = layer._forward_bf(st, memory, mem.x_lp, mempos, mem.xpos_lp, query_pos, tgt_mask_kp, mem_mask_kp,
                                   last or self.track_attention)
            if self.return_intermediate:
                intermediate.append(st.x)
        output = torch.stack(intermediate) if self.return_intermediate else st.x
        if self.norm is not None:                                  # :322-325
            return add_layer_norm(output, None, self.norm)[0], output
        return output, output

    def forward(self, tgt, memory, tgt_mask=None, memory_mask=None, tgt_key_padding_mask=None,
                memory_key_padding_mask=None, pos=None, query_pos=None, prev_frame=None):
        _check_no_attn_mask(tgt_mask, memory_mask)
        hs, raw = self._forward_bf(_bf(tgt), _bf(memory), tgt_key_padding_mask, memory_key_padding_mask, _bf(pos),
                                   _bf(query_pos))
        return hs.transpose(-3, -2), raw.transpose(-3, -2)


class KinematicTransformer(nn.Module):
    def __init__(self, d_model=512, nhead=8, num_encoder_layers=6, num_decoder_layers=6, dim_feedforward=2048,
                 dropout=0.1, activation="relu", normalize_before=False, return_intermediate_dec=False,
                 track_attention=False):
        super().__init__()
        self.d_model = d_model
        self.nhead = nhead
        encoder_layer = TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, activation, normalize_before)
        encoder_norm = nn.LayerNorm(d_model) if normalize_before else None
        self.encoder = TransformerEncoder(encoder_layer, num_encoder_layers, encoder_norm)
        decoder_layer = TransformerDecoderLayer(d_model, nhead, dim_feedforward, dropout, activation, normalize_before)
        decoder_norm = nn.LayerNorm(d_model)
        self.decoder = TransformerDecoder(decoder_layer, encoder_layer, num_decoder_layers, decoder_norm,
                                          return_intermediate=return_intermediate_dec, track_attention=track_attention)
        self._reset_parameters()

    def _reset_parameters(self):
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)

    def _forward_bf(self, src, mask, query_embed, tgt, pos_src, query_padding_mask=None):
        """src / pos_src (BS, T, C); query_embed / tgt (BS, Lq, C).  Batch-first (hs, hs_without_norm, memory).
        query_padding_mask (BS, Lq) bool: queries that are padding (a caller that pads the query list to a bucket, e.g.
        KineFrameGraphs) -- masked as KEYS of the decoder's self-attention, so the real queries see exactly what they
        would see unpadded (the reference passes tgt_key_padding_mask=None: it never pads queries)."""
        if tgt is None:
            tgt = torch.zeros_like(query_embed)
        memory = self.encoder._forward_bf(src, mask, pos_src)
        hs, raw = self.decoder._forward_bf(tgt, memory, query_padding_mask, mask, pos_src, query_embed)
        return hs, raw, memory

    def forward(self, src, mask, query_embed, tgt=None, pos_src=None):
        """Reference layouts (:122-142): src / pos_src (BS, T, C); query_embed / tgt (Lq, BS, C); returns the
        sequence-first (hs, hs_without_norm, memory)."""
        hs, raw, memory = self._forward_bf(src.contiguous(), mask, _bf(query_embed), _bf(tgt),
                                           None if pos_src is None else pos_src.contiguous())
        return hs.transpose(-3, -2), raw.transpose(-3, -2), memory.transpose(0, 1)


class IntertwinedBranch(nn.Module):
    def __init__(self, d_model=256, dropout=0.1, activation="relu", dim_concat=3):
        super().__init__()
        self.linear_input1 = nn.Linear(d_model, d_model // 2)
        self.linear_input2 = nn.Linear(d_model, d_model // 2)
        self.dropout = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_model // 2, d_model)        # (unused by the reference's forward too, :478)
        self.norm = nn.LayerNorm(d_model)
        self.activation = _get_activation_fn(activation)
        self.dim_concat = dim_concat

    def forward(self, src1, src2):
        """norm(dropout(activation(cat[linear_input1(src1), linear_input2(src2)])) + src1)   (:483-487); the
        concatenation runs over the channel dimension (dim_concat = the last dimension of the inputs)."""
        if self.dim_concat != src1.dim() - 1:
            raise RuntimeError("kinet_b200: IntertwinedBranch concatenates over the channel dimension")
        lp = _lp_dtype(src1)
        a = src1 if src1.dtype == lp else src1.to(lp)
        b = src2 if src2.dtype == lp else src2.to(lp)
        x = torch.cat([_act_linear(a, self.linear_input1, self.activation),
                       _act_linear(b, self.linear_input2, self.activation)], -1)
        return add_layer_norm(src1.contiguous(), self.dropout(x), self.norm)[0]


class DualKinematicTransformer(nn.Module):
    def __init__(self, d_model=512, nhead=8, num_encoder_layers=6, num_decoder_layers=6, dim_feedforward=2048,
                 dropout=0.1, activation="relu", normalize_before=False, return_intermediate_dec=False,
                 track_attention=False):
        super().__init__()
        self.d_model = d_model
        args = (d_model, nhead, num_encoder_layers, num_decoder_layers, dim_feedforward, dropout, activation,
                normalize_before, return_intermediate_dec, track_attention)
        self.transformer_det = KinematicTransformer(*args)
        self.transformer_metadata = KinematicTransformer(*args)
        self.detection_branch = IntertwinedBranch(d_model, dropout, activation)
        self.metadata_branch = IntertwinedBranch(d_model, dropout, activation)

    def forward(self, src_boxes, src_metadata, mask, query_embed_bbox, query_embed_metadata, tgt_bboxes, tgt_metadata,
                pos_boxes=None, pos_metadata=None):
        """Reference layouts (:170-185): src_* / pos_* (BS, T, C), mask (BS, T) bool, query_embed_* / tgt_*
        (Lq, BS, C).  Returns (hs_det, hs_metadata, hs_without_norm_det) as (layers, BS, Lq, C) and memory_det
        (BS, T, C) -- the reference's ``transpose(1, 2)`` / ``permute(1, 0, 2)`` of its sequence-first tensors."""
        return self.forward_masked_queries(src_boxes, src_metadata, mask, query_embed_bbox, query_embed_metadata,
                                           tgt_bboxes, tgt_metadata, pos_boxes, pos_metadata, None)

    def forward_masked_queries(self, src_boxes, src_metadata, mask, query_embed_bbox, query_embed_metadata, tgt_bboxes,
                               tgt_metadata, pos_boxes=None, pos_metadata=None, query_padding_mask=None):
        """forward() with a (BS, Lq) bool mask of PADDING queries (see KinematicTransformer._forward_bf)."""
        c = lambda t: None if t is None else t.contiguous()
        (hs_det, raw_det, memory_det), (hs_meta, _, _) = _two_branches(
            lambda: self.transformer_det._forward_bf(c(src_boxes), mask, _bf(query_embed_bbox), _bf(tgt_bboxes),
                                                     c(pos_boxes), query_padding_mask),
            lambda: self.transformer_metadata._forward_bf(c(src_metadata), mask, _bf(query_embed_metadata),
                                                          _bf(tgt_metadata), c(pos_metadata), query_padding_mask),
            src_boxes.device)
        # the concatenation dimension of the reference (3: channels of a (layers, Lq, BS, C) stack) stays the last one
        self.detection_branch.dim_concat = self.metadata_branch.dim_concat = hs_det.dim() - 1
        hs_det = self.detection_branch(hs_det, hs_meta)
        hs_meta = self.metadata_branch(hs_meta, hs_det)
        return hs_det, hs_meta, raw_det, memory_det


class DualKinematicEncoder(nn.Module):
    def __init__(self, d_model=512, nhead=8, num_encoder_layers=6, dim_feedforward=2048, dropout=0.1,
                 activation="relu", normalize_before=False):
        super().__init__()
        self.d_model = d_model
        encoder_layer_det = TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, activation, normalize_before)
        encoder_norm_det = nn.LayerNorm(d_model) if normalize_before else None
        self.encoder_det = TransformerEncoder(encoder_layer_det, num_encoder_layers, encoder_norm_det)
        encoder_layer_meta = TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, activation, normalize_before)
        encoder_norm_meta = nn.LayerNorm(d_model) if normalize_before else None
        self.encoder_meta = TransformerEncoder(encoder_layer_meta, num_encoder_layers, encoder_norm_meta)
        self.detection_branch = IntertwinedBranch(d_model, dropout, activation, dim_concat=2)
        self.metadata_branch = IntertwinedBranch(d_model, dropout, activation, dim_concat=2)

    def forward(self, src_boxes, src_metadata, mask, pos_boxes=None, pos_metadata=None):
        """(:217-242) returns (hs_det[None], hs_metadata[None]) as (1, BS, T, C) and the two memories as (BS, C, T)."""
        c = lambda t: None if t is None else t.contiguous()
        memory_det, memory_meta = _two_branches(
            lambda: self.encoder_det._forward_bf(c(src_boxes), mask, c(pos_boxes)),
            lambda: self.encoder_meta._forward_bf(c(src_metadata), mask, c(pos_metadata)), src_boxes.device)
        hs_det = self.detection_branch(memory_det, memory_meta)
        hs_meta = self.metadata_branch(memory_meta, memory_det)
        return hs_det[None], hs_meta[None], memory_meta.transpose(1, 2), memory_det.transpose(1, 2)


def build_transformer(args):
    """The `kine` branch of the reference's build_transformer (:503-525)."""
    if not getattr(args, "kine", False):
        raise RuntimeError("kinet_b200.modules.kinematic_transformer builds the kine transformers only")
    if getattr(args, "use_encoder_only", False):
        return DualKinematicEncoder(d_model=args.hidden_dim, nhead=args.nheads, num_encoder_layers=args.enc_layers,
                                    dim_feedforward=args.dim_feedforward, dropout=args.dropout,
                                    activation=args.activation, normalize_before=args.pre_norm)
    return DualKinematicTransformer(d_model=args.hidden_dim, nhead=args.nheads, num_encoder_layers=args.enc_layers,
                                    num_decoder_layers=args.dec_layers, dim_feedforward=args.dim_feedforward,
                                    dropout=args.dropout, activation=args.activation, normalize_before=args.pre_norm,
                                    return_intermediate_dec=True, track_attention=args.track_attention)
