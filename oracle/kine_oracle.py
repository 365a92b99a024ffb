"""CPU restatement of the reference's kinematic (`kine`) transformers.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): only tests/ may import it.

Plain torch, any float dtype, written against /root/reference/src/trackformer/models/transformer.py:

  mha                    what nn.MultiheadAttention computes for the calls at :331-333, :393-401: in_proj, per-head
                         softmax(q k^T / sqrt(d) + key_padding_mask) v, out_proj (no attention dropout: eval / p = 0)
  encoder_layer          TransformerEncoderLayer.forward_post :323-339 / forward_pre :341-354
  decoder_layer          TransformerDecoderLayer.forward_post :385-407 / forward_pre :409-431
  encoder / decoder      TransformerEncoder :245-266, TransformerDecoder :269-326 (without track_attention)
  kinematic_transformer  KinematicTransformer.forward :122-142
  intertwined_branch     IntertwinedBranch.forward :483-487
  dual_kinematic_transformer / dual_kinematic_encoder    :170-185 / :217-242

Tensors are sequence-first (T, BS, C) inside, as in the reference.  Parameters are passed as a flat dict
name -> tensor with the reference's state_dict names, so the golden file's parameters (tests/golden/kine_f32.npz,
produced by the reference's own classes) plug in directly.

Parity pin: tests/test_kine.py::test_kine_oracle_matches_reference_golden checks outputs and gradients against the
golden file.
"""
import torch
import torch.nn.functional as F


def _lin(p, name, x):
    return F.linear(x, p[name + ".weight"], p[name + ".bias"])


def _ln(p, name, x):
    return F.layer_norm(x, (x.shape[-1],), p[name + ".weight"], p[name + ".bias"], 1e-5)


def mha(p, pre, q_in, k_in, v_in, key_padding_mask, nhead):
    Lq, N, C = q_in.shape
    Lk = k_in.shape[0]
    d = C // nhead
    w, b = p[pre + ".in_proj_weight"], p[pre + ".in_proj_bias"]
    q = F.linear(q_in, w[:C], b[:C]).reshape(Lq, N, nhead, d).permute(1, 2, 0, 3)
    k = F.linear(k_in, w[C:2 * C], b[C:2 * C]).reshape(Lk, N, nhead, d).permute(1, 2, 0, 3)
    v = F.linear(v_in, w[2 * C:], b[2 * C:]).reshape(Lk, N, nhead, d).permute(1, 2, 0, 3)
    s = (q @ k.transpose(-1, -2)) * d ** -0.5
    if key_padding_mask is not None:
        s = s.masked_fill(key_padding_mask[:, None, None, :], float("-inf"))
    o = (s.softmax(-1) @ v).permute(2, 0, 1, 3).reshape(Lq, N, C)
    return _lin(p, pre + ".out_proj", o)


def _pe(x, pos):
    return x if pos is None else x + pos


def encoder_layer(p, pre, src, mask, pos, nhead, pre_norm):
    if not pre_norm:
        q = _pe(src, pos)
        src = _ln(p, pre + "norm1", src + mha(p, pre + "self_attn", q, q, src, mask, nhead))
        return _ln(p, pre + "norm2", src + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", src))))
    s2 = _ln(p, pre + "norm1", src)
    q = _pe(s2, pos)
    src = src + mha(p, pre + "self_attn", q, q, s2, mask, nhead)
    s2 = _ln(p, pre + "norm2", src)
    return src + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", s2)))


def decoder_layer(p, pre, tgt, memory, mem_mask, pos, query_pos, nhead, pre_norm):
    if not pre_norm:
        q = _pe(tgt, query_pos)
        tgt = _ln(p, pre + "norm1", tgt + mha(p, pre + "self_attn", q, q, tgt, None, nhead))
        tgt = _ln(p, pre + "norm2", tgt + mha(p, pre + "multihead_attn", _pe(tgt, query_pos), _pe(memory, pos), memory,
                                                mem_mask, nhead))
        return _ln(p, pre + "norm3", tgt + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", tgt))))
    t2 = _ln(p, pre + "norm1", tgt)
    q = _pe(t2, query_pos)
    tgt = tgt + mha(p, pre + "self_attn", q, q, t2, None, nhead)
    t2 = _ln(p, pre + "norm2", tgt)
    tgt = tgt + mha(p, pre + "multihead_attn", _pe(t2, query_pos), _pe(memory, pos), memory, mem_mask, nhead)
    t2 = _ln(p, pre + "norm3", tgt)
    return tgt + _lin(p, pre + "linear2", F.relu(_lin(p, pre + "linear1", t2)))


def encoder(p, pre, src, mask, pos, nhead, layers, pre_norm):
    for i in range(layers):
        src = encoder_layer(p, "%slayers.%d." % (pre, i), src, mask, pos, nhead, pre_norm)
    return _ln(p, pre + "norm", src) if pre_norm else src


def decoder(p, pre, tgt, memory, mem_mask, pos, query_pos, nhead, layers, pre_norm, return_intermediate):
    inter = []
    for i in range(layers):
        tgt = decoder_layer(p, "%slayers.%d." % (pre, i), tgt, memory, mem_mask, pos, query_pos, nhead, pre_norm)
        inter.append(tgt)
    out = torch.stack(inter) if return_intermediate else tgt
    return _ln(p, pre + "norm", out), out


def kinematic_transformer(p, pre, src, mask, query_embed, tgt, pos_src, nhead, enc_layers, dec_layers, pre_norm=False,
                          return_intermediate=True):
    """src / pos_src (BS, T, C); query_embed / tgt (Lq, BS, C) -> sequence-first (hs, hs_without_norm, memory)."""
    src = src.permute(1, 0, 2)
    pos = None if pos_src is None else pos_src.permute(1, 0, 2)
    if tgt is None:
        tgt = torch.zeros_like(query_embed)
    memory = encoder(p, pre + "encoder.", src, mask, pos, nhead, enc_layers, pre_norm)
    hs, raw = decoder(p, pre + "decoder.", tgt, memory, mask, pos, query_embed, nhead, dec_layers, pre_norm,
                      return_intermediate)
    return hs, raw, memory


def intertwined_branch(p, pre, src1, src2):
    x = F.relu(torch.cat([_lin(p, pre + "linear_input1", src1), _lin(p, pre + "linear_input2", src2)], -1))
    return _ln(p, pre + "norm", x + src1)


def dual_kinematic_transformer(p, src_b, src_m, mask, qe_b, qe_m, tgt_b, tgt_m, pos_b, pos_m, nhead, enc_layers,
                               dec_layers):
    hs_d, raw_d, mem_d = kinematic_transformer(p, "transformer_det.", src_b, mask, qe_b, tgt_b, pos_b, nhead,
                                               enc_layers, dec_layers)
    hs_m, _, _ = kinematic_transformer(p, "transformer_metadata.", src_m, mask, qe_m, tgt_m, pos_m, nhead,
                                       enc_layers, dec_layers)
    hs_d = intertwined_branch(p, "detection_branch.", hs_d, hs_m)
    hs_m = intertwined_branch(p, "metadata_branch.", hs_m, hs_d)
    return hs_d.transpose(1, 2), hs_m.transpose(1, 2), raw_d.transpose(1, 2), mem_d.permute(1, 0, 2)


def dual_kinematic_encoder(p, src_b, src_m, mask, pos_b, pos_m, nhead, layers):
    mem_d = encoder(p, "encoder_det.", src_b.permute(1, 0, 2), mask, pos_b.permute(1, 0, 2), nhead, layers, False)
    mem_m = encoder(p, "encoder_meta.", src_m.permute(1, 0, 2), mask, pos_m.permute(1, 0, 2), nhead, layers, False)
    hs_d = intertwined_branch(p, "detection_branch.", mem_d, mem_m)
    hs_m = intertwined_branch(p, "metadata_branch.", mem_m, mem_d)
    return hs_d.transpose(0, 1)[None], hs_m.transpose(0, 1)[None], mem_m.permute(1, 2, 0), mem_d.permute(1, 2, 0)
