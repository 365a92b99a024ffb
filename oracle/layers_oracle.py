"""CPU restatement of the reference's deformable-transformer layers around MSDeformAttn.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): only tests/ may import it.

Plain torch on CPU, any float dtype, written against the reference's source (all paths relative to
/root/reference/src/trackformer/models/):

  msdeform_attn_module   ops/modules/ms_deform_attn.py:52-89   (incl. the (H, W) offset normalisation :77-79
                                                               and query_attn_mask :73-74)
  encoder_layer          deformable_transformer.py:289-299 (+ forward_ffn :282-287)
  encoder_reference_points / encoder        :308-331
  decoder_layer          :363-386 (self-attention as nn.MultiheadAttention computes it: in_proj, per-head
                         softmax(q k^T / sqrt(d) + key_padding_mask) v, out_proj)
  decoder                :400-433 (without box refinement: bbox_embed is None)

Parameters are passed as a flat dict name -> tensor with the reference's state_dict names, so the golden
file's parameters (tests/golden/layers_f32.npz, produced by the reference's own classes) plug in directly.
The sampling core is oracle.msda_oracle.torch_port_forward (the port of ms_deform_attn_core_pytorch).
Dropout is the identity (eval mode / p = 0).

Parity pin: tests/test_oracle.py::test_layers_oracle_matches_reference_golden checks outputs and gradients
against the golden file.
"""
import torch
import torch.nn.functional as F

from .msda_oracle import torch_port_forward


def _lin(p, name, x):
    return F.linear(x, p[name + ".weight"], p[name + ".bias"])


def _ln(p, name, x):
    return F.layer_norm(x, (x.shape[-1],), p[name + ".weight"], p[name + ".bias"], 1e-5)


def msdeform_attn_module(p, pre, query, reference_points, input_flatten, shapes, padding_mask, query_mask,
                         n_heads, n_levels, n_points):
    N, Lq, C = query.shape
    S = input_flatten.shape[1]
    value = _lin(p, pre + "value_proj", input_flatten)
    if padding_mask is not None:
        value = value.masked_fill(padding_mask[..., None], 0.0)
    value = value.view(N, S, n_heads, C // n_heads)
    off = _lin(p, pre + "sampling_offsets", query).view(N, Lq, n_heads, n_levels, n_points, 2)
    attn = _lin(p, pre + "attention_weights", query).view(N, Lq, n_heads, n_levels * n_points)
    attn = F.softmax(attn, -1).view(N, Lq, n_heads, n_levels, n_points)
    if query_mask is not None:
        attn = attn.masked_fill(query_mask[..., None, None, None], 0.0)
    if reference_points.shape[-1] == 2:
        # the reference divides by input_spatial_shapes as stored, i.e. (H, W): x / H, y / W (:77-79)
        loc = reference_points[:, :, None, :, None, :] + off / shapes[None, None, None, :, None, :].to(off.dtype)
    else:
        loc = reference_points[:, :, None, :, None, :2] + off / n_points * reference_points[:, :, None, :, None, 2:] * 0.5
    out = torch_port_forward(value, shapes, loc, attn)
    return _lin(p, pre + "output_proj", out)


def encoder_layer(p, pre, src, pos, reference_points, shapes, padding_mask, cfg):
    q = src if pos is None else src + pos
    src = _ln(p, pre + "norm1", src + msdeform_attn_module(p, pre + "self_attn.", q, reference_points, src, shapes,
                                                          padding_mask, None, *cfg))
    return _ln(p, pre + "norm2", src + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", src))))


def encoder_reference_points(levels, valid_ratios):
    refs = []
    for lvl, (H, W) in enumerate(levels):
        ys = torch.linspace(0.5, H - 0.5, H, dtype=valid_ratios.dtype)
        xs = torch.linspace(0.5, W - 0.5, W, dtype=valid_ratios.dtype)
        gy, gx = torch.meshgrid(ys, xs, indexing="ij")
        ry = gy.reshape(-1)[None] / (valid_ratios[:, None, lvl, 1] * H)
        rx = gx.reshape(-1)[None] / (valid_ratios[:, None, lvl, 0] * W)
        refs.append(torch.stack((rx, ry), -1))
    return torch.cat(refs, 1)[:, :, None] * valid_ratios[:, None]


def encoder(p, pre, n_layers, src, levels, shapes, valid_ratios, pos, padding_mask, cfg):
    ref = encoder_reference_points(levels, valid_ratios)
    for i in range(n_layers):
        src = encoder_layer(p, "%slayers.%d." % (pre, i), src, pos, ref, shapes, padding_mask, cfg)
    return src


def multihead_self_attention(p, pre, q_in, v_in, key_padding_mask, n_heads):
    """nn.MultiheadAttention(q, k = q, v) on (N, Lq, C) inputs (batch first here; the reference transposes)."""
    N, Lq, C = q_in.shape
    w, b = p[pre + "in_proj_weight"], p[pre + "in_proj_bias"]
    q = F.linear(q_in, w[:C], b[:C]).view(N, Lq, n_heads, C // n_heads).transpose(1, 2)
    k = F.linear(q_in, w[C:2 * C], b[C:2 * C]).view(N, Lq, n_heads, C // n_heads).transpose(1, 2)
    v = F.linear(v_in, w[2 * C:], b[2 * C:]).view(N, Lq, n_heads, C // n_heads).transpose(1, 2)
    logits = q @ k.transpose(-1, -2) / (C // n_heads) ** 0.5
    if key_padding_mask is not None:
        logits = logits.masked_fill(key_padding_mask[:, None, None, :], float("-inf"))
    out = (F.softmax(logits, -1) @ v).transpose(1, 2).reshape(N, Lq, C)
    return F.linear(out, p[pre + "out_proj.weight"], p[pre + "out_proj.bias"])


def decoder_layer(p, pre, tgt, query_pos, reference_points, memory, shapes, padding_mask, query_mask, cfg):
    q = tgt if query_pos is None else tgt + query_pos
    tgt = _ln(p, pre + "norm2", tgt + multihead_self_attention(p, pre + "self_attn.", q, tgt, query_mask, cfg[0]))
    q = tgt if query_pos is None else tgt + query_pos
    tgt = _ln(p, pre + "norm1", tgt + msdeform_attn_module(p, pre + "cross_attn.", q, reference_points, memory, shapes,
                                                          padding_mask, query_mask, *cfg))
    return _ln(p, pre + "norm3", tgt + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", tgt))))


def decoder(p, pre, n_layers, tgt, reference_points, memory, shapes, valid_ratios, query_pos, padding_mask,
            query_mask, cfg):
    for i in range(n_layers):
        if reference_points.shape[-1] == 4:
            rin = reference_points[:, :, None] * torch.cat([valid_ratios, valid_ratios], -1)[:, None]
        else:
            rin = reference_points[:, :, None] * valid_ratios[:, None]
        tgt = decoder_layer(p, "%slayers.%d." % (pre, i), tgt, query_pos, rin, memory, shapes, padding_mask,
                            query_mask, cfg)
    return tgt, reference_points
