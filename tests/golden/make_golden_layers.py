#!/usr/bin/env python
"""Generate tests/golden/layers_f32.npz from the REAL reference encoder / decoder classes.

Run in the build container only (needs /root/reference; the GPU box has none):

    python tests/golden/make_golden_layers.py

Imports the reference's src/trackformer/models/deformable_transformer.py unmodified
(DeformableTransformerEncoder(Layer) :260-331, DeformableTransformerDecoder(Layer) :334-433) with its own
MSDeformAttn module (ops/modules/ms_deform_attn.py).  The module's CUDA-only ``MSDeformAttnFunction.apply``
is pointed at the reference's own CPU-capable ``ms_deform_attn_core_pytorch``
(ops/functions/ms_deform_attn_func.py:34-54) -- the substitution SURVEY.md section 8(c) names.  Package
``__init__`` files are bypassed (they import the whole project); ``visdom`` and the CUDA extension are
empty stubs that the imported code never calls.

Stored: every parameter of a 2-layer encoder + 2-layer decoder (d_model 128, 8 heads, 2 levels, 2 points),
the inputs, the encoder memory, the decoder output and reference points, and the gradients of a fixed loss
with respect to the inputs, every 1-d parameter and the weight matrices of layer 0 (fp32, CPU, dropout off).
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

SRC = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference():
    def pkg(name, path):
        m = types.ModuleType(name)
        m.__path__ = [path]
        sys.modules[name] = m
    pkg("trackformer", SRC + "/trackformer")
    pkg("trackformer.models", SRC + "/trackformer/models")
    pkg("trackformer.util", SRC + "/trackformer/util")
    vis = types.ModuleType("visdom")
    vis.Visdom = object
    sys.modules["visdom"] = vis
    sys.modules["MultiScaleDeformableAttention"] = types.ModuleType("MultiScaleDeformableAttention")
    dt = importlib.import_module("trackformer.models.deformable_transformer")
    mod = importlib.import_module("trackformer.models.ops.modules.ms_deform_attn")
    fn = importlib.import_module("trackformer.models.ops.functions.ms_deform_attn_func")

    class CoreFunction:
        @staticmethod
        def apply(value, shapes, loc, attn, im2col_step):
            return fn.ms_deform_attn_core_pytorch(value, shapes, loc, attn)
    mod.MSDeformAttnFunction = CoreFunction
    return dt


def main():
    dt = load_reference()
    torch.manual_seed(7)
    d, ffn, L, M, P, N, Lq = 128, 64, 2, 8, 2, 2, 5
    levels = [(6, 8), (3, 4)]
    S = sum(h * w for h, w in levels)
    enc = dt.DeformableTransformerEncoder(dt.DeformableTransformerEncoderLayer(d, ffn, 0.0, "relu", L, M, P), 2)
    dec = dt.DeformableTransformerDecoder(dt.DeformableTransformerDecoderLayer(d, ffn, 0.0, "relu", L, M, P), 2)
    # non-degenerate parameters everywhere (the module zero-initialises offsets / attention weights; clones share values)
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():
        for name, p in list(enc.named_parameters()) + list(dec.named_parameters()):
            if "sampling_offsets.bias" in name:
                p.add_(torch.randn(p.shape, generator=g) * 0.1)
            elif name.endswith("norm1.weight") or name.endswith("norm2.weight") or name.endswith("norm3.weight"):
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=g))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (0.5 / p.shape[1] ** 0.5))
    enc.eval(); dec.eval()

    shapes = torch.as_tensor(levels, dtype=torch.long)
    src = torch.randn(N, S, d, generator=g).requires_grad_(True)
    pos = (0.3 * torch.randn(N, S, d, generator=g)).requires_grad_(True)
    valid = 0.8 + 0.2 * torch.rand(N, L, 2, generator=g)
    mask = torch.zeros(N, S, dtype=torch.bool)
    mask[1, -3:] = True                                    # a few padded positions in the coarse level
    tgt = torch.randn(N, Lq, d, generator=g).requires_grad_(True)
    qpos = (0.3 * torch.randn(N, Lq, d, generator=g)).requires_grad_(True)
    ref = torch.rand(N, Lq, 2, generator=g) * 0.8 + 0.1
    qmask = torch.zeros(N, Lq, dtype=torch.bool)
    qmask[0, -1] = True                                    # one masked (padding) query
    w_mem = torch.randn(N, S, d, generator=g)
    w_out = torch.randn(N, Lq, d, generator=g)

    memory = enc(src, shapes, valid, pos, mask)
    out, ref_out = dec(tgt, ref, memory, shapes, valid, qpos, mask, qmask)
    loss = (memory * w_mem).sum() + (out * w_out).sum()
    params = dict(("enc." + k, v) for k, v in enc.named_parameters())
    params.update(("dec." + k, v) for k, v in dec.named_parameters())
    names = list(params)
    grads = torch.autograd.grad(loss, [src, pos, tgt, qpos] + [params[k] for k in names])

    blob = dict(levels=np.asarray(levels), src=src.detach().numpy(), pos=pos.detach().numpy(), valid=valid.numpy(),
                mask=mask.numpy(), tgt=tgt.detach().numpy(), qpos=qpos.detach().numpy(), ref=ref.numpy(),
                qmask=qmask.numpy(), w_mem=w_mem.numpy(), w_out=w_out.numpy(),
                memory=memory.detach().numpy(), out=out.detach().numpy(), ref_out=ref_out.numpy(),
                g_src=grads[0].numpy(), g_pos=grads[1].numpy(), g_tgt=grads[2].numpy(), g_qpos=grads[3].numpy())
    for k, gk in zip(names, grads[4:]):
        blob["p:" + k] = params[k].detach().numpy()
        if gk.dim() == 1 or ".layers.0." in k:      # every bias / norm gradient; weight-matrix gradients of layer 0
            blob["g:" + k] = gk.numpy()
    path = os.path.join(HERE, "layers_f32.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes;", len(names), "parameters")


if __name__ == "__main__":
    main()
