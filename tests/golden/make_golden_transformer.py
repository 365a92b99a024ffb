#!/usr/bin/env python
"""Generate tests/golden/transformer_f32.npz from the REAL reference DeformableTransformer.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_transformer.py

Same import recipe as make_golden_layers.py (reference file unmodified, package __init__ bypassed, the module's
CUDA-only MSDeformAttnFunction.apply pointed at the reference's own ms_deform_attn_core_pytorch).  Two cases of
src/trackformer/models/deformable_transformer.py:19-257 (fp32, CPU, dropout 0, return_intermediate_dec=True):

  a: d_model 128, 2 + 2 layers, 2 feature levels, 7 object queries + 3 track queries handed in through
     ``targets`` ('track_query_hs_embeds' / 'track_query_boxes', :203-229); outputs and input gradients
  b: d_model 64, 1 + 1 layers, 4 feature levels with multi_frame_attention_separate_encoder=True (the encoder
     runs once per frame on 2 levels each, the decoder sees all 4, :160-175); outputs only
  c: d_model 256 (parameters are a function of their names, not stored), 1 + 2 layers, two_stage=True with 6 proposals (:181-196) and iterative box refinement in the
     decoder (bbox_embed set, :413-424; 4-d reference points); outputs only
"""
import os

import numpy as np
import torch

from make_golden_layers import load_reference

HERE = os.path.dirname(os.path.abspath(__file__))


def randomise(model, g):
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "sampling_offsets.bias" in name:
                p.add_(torch.randn(p.shape, generator=g) * 0.1)
            elif "norm" in name and name.endswith("weight"):
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=g))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (0.5 / p.shape[-1] ** 0.5))


def randomise_by_name(model):
    """Parameters as a pure function of their NAME (seed = crc32 of the name): a test can rebuild them without the
    golden file having to carry a d_model = 256 model (two-stage hard-codes 4 x 128 sine features, :71)."""
    import zlib
    with torch.no_grad():
        for name, p in model.named_parameters():
            gn = torch.Generator().manual_seed(zlib.crc32(name.encode()))
            if "norm" in name and name.endswith("weight"):
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=gn))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=gn))
            else:
                p.copy_(torch.randn(p.shape, generator=gn) * (0.5 / p.shape[-1] ** 0.5))


def inputs(g, d, sizes, N):
    srcs = [torch.randn(N, d, h, w, generator=g).requires_grad_(True) for h, w in sizes]
    poss = [0.3 * torch.randn(N, d, h, w, generator=g) for h, w in sizes]
    masks = []
    for h, w in sizes:
        m = torch.zeros(N, h, w, dtype=torch.bool)
        m[1, :, w - max(1, w // 4):] = True              # second image is narrower: right-hand padding
        masks.append(m)
    return srcs, masks, poss


def main():
    dt = load_reference()
    blob = {}
    g = torch.Generator().manual_seed(21)

    # ---- case a ----
    d, N, Lq, T = 128, 2, 7, 3
    sizes = [(6, 8), (3, 4)]
    model = dt.DeformableTransformer(d_model=d, nhead=8, num_encoder_layers=2, num_decoder_layers=2, dim_feedforward=64,
                                     dropout=0.0, activation="relu", return_intermediate_dec=True,
                                     num_feature_levels=2, dec_n_points=2, enc_n_points=2)
    randomise(model, g)
    model.eval()
    srcs, masks, poss = inputs(g, d, sizes, N)
    query_embed = torch.randn(Lq, 2 * d, generator=g).requires_grad_(True)
    targets = [dict(track_query_hs_embeds=torch.randn(T, d, generator=g), track_query_boxes=torch.rand(T, 4, generator=g) * 0.6 + 0.2)
               for _ in range(N)]
    hs, memory, init_ref, inter_ref, _, _ = model(srcs, masks, poss, query_embed, targets)
    w_hs, w_mem = torch.randn(hs.shape, generator=g), torch.randn(memory.shape, generator=g)
    grads = torch.autograd.grad((hs * w_hs).sum() + (memory * w_mem).sum(), srcs + [query_embed])
    for k, v in model.state_dict().items():
        blob["a.p:" + k] = v.numpy()
    for i, (s, m, p) in enumerate(zip(srcs, masks, poss)):
        blob["a.src%d" % i], blob["a.mask%d" % i], blob["a.pos%d" % i] = s.detach().numpy(), m.numpy(), p.numpy()
        blob["a.g_src%d" % i] = grads[i].numpy()
    blob.update({"a.query_embed": query_embed.detach().numpy(), "a.g_query_embed": grads[-1].numpy(),
                 "a.track_hs": torch.stack([t["track_query_hs_embeds"] for t in targets]).numpy(),
                 "a.track_boxes": torch.stack([t["track_query_boxes"] for t in targets]).numpy(),
                 "a.hs": hs.detach().numpy(), "a.memory": memory.detach().numpy(), "a.init_ref": init_ref.detach().numpy(),
                 "a.inter_ref": inter_ref.detach().numpy(), "a.w_hs": w_hs.numpy(), "a.w_mem": w_mem.numpy()})

    # ---- case b ----
    d, N, Lq = 64, 2, 5
    sizes = [(6, 8), (3, 4), (6, 8), (3, 4)]              # current frame's levels, then the previous frame's
    model = dt.DeformableTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1, dim_feedforward=32,
                                     dropout=0.0, activation="relu", return_intermediate_dec=True,
                                     num_feature_levels=4, dec_n_points=2, enc_n_points=2,
                                     multi_frame_attention_separate_encoder=True)
    randomise(model, g)
    model.eval()
    srcs, masks, poss = inputs(g, d, sizes, N)
    query_embed = torch.randn(Lq, 2 * d, generator=g)
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, _, _ = model(srcs, masks, poss, query_embed, None)
    for k, v in model.state_dict().items():
        blob["b.p:" + k] = v.numpy()
    for i, (s, m, p) in enumerate(zip(srcs, masks, poss)):
        blob["b.src%d" % i], blob["b.mask%d" % i], blob["b.pos%d" % i] = s.detach().numpy(), m.numpy(), p.numpy()
    blob.update({"b.query_embed": query_embed.numpy(), "b.hs": hs.numpy(), "b.memory": memory.numpy(),
                 "b.init_ref": init_ref.numpy(), "b.inter_ref": inter_ref.numpy()})

    # ---- case c: two-stage (proposals from the encoder output) + iterative box refinement in the decoder ----
    d, N = 256, 2
    sizes = [(6, 8), (3, 4)]
    model = dt.DeformableTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2, dim_feedforward=32,
                                     dropout=0.0, activation="relu", return_intermediate_dec=True,
                                     num_feature_levels=2, dec_n_points=2, enc_n_points=2,
                                     two_stage=True, two_stage_num_proposals=6)
    # the detection head normally installs these (deformable_detr.py); plain Linear layers stand in for them
    model.decoder.class_embed = torch.nn.ModuleList(torch.nn.Linear(d, 3) for _ in range(3))
    model.decoder.bbox_embed = torch.nn.ModuleList(torch.nn.Linear(d, 4) for _ in range(3))
    randomise_by_name(model)
    model.eval()
    srcs, masks, poss = inputs(g, d, sizes, N)
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, enc_cls, enc_box = model(srcs, masks, poss, None, None)
    for i, (s_, m, p) in enumerate(zip(srcs, masks, poss)):
        blob["c.src%d" % i], blob["c.mask%d" % i], blob["c.pos%d" % i] = s_.detach().numpy(), m.numpy(), p.numpy()
    blob.update({"c.hs": hs.numpy(), "c.memory": memory.numpy(), "c.init_ref": init_ref.numpy(),
                 "c.inter_ref": inter_ref.numpy(), "c.enc_cls": enc_cls.numpy(), "c.enc_box": enc_box.numpy()})

    path = os.path.join(HERE, "transformer_f32.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
