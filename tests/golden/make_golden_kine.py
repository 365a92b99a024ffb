#!/usr/bin/env python
"""Generate tests/golden/kine_f32.npz from the REAL reference kinematic transformers.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_kine.py

Imports the reference's src/trackformer/models/transformer.py unmodified (it depends on torch only) and runs, in
fp32 on the CPU with dropout 0:

  a: DualKinematicTransformer (:145-185) d_model 72 (8 heads of 9), 2 + 2 layers, ffn 96, return_intermediate_dec=True;
     BS 2, 11 detection tokens (3 of the second sample padded), 7 queries of which the first 2 carry a non-zero
     ``tgt`` (track queries, detr.py:377-401); parameters stored; outputs, input gradients, gradients of every 1-d
     parameter and of the weight matrices of the first layers
  b: the same class at the kinet configuration's width (cfgs/train_kinet.yaml:17-27: d_model 288 = 8 heads of 36,
     1 + 1 layers, 150 queries) over 150 + 10 tokens, ffn 64; parameters are a function of their names
     (make_golden_transformer.randomise_by_name), not stored; outputs only
  c: DualKinematicEncoder (:188-242), d_model 72, 2 layers; name-seeded parameters; outputs only
  d: KinematicTransformer (:85-142) with normalize_before=True (pre-norm) and return_intermediate_dec=False, d_model 72,
     1 + 2 layers; name-seeded parameters; outputs only
  e: KinematicDetectorTransformer (src/trackformer/models/detr.py:288-417; imported with the package __init__ files
     bypassed, as make_golden_layers.py does) around a d_model 72, 1 + 2 layer DualKinematicTransformer, 6 queries,
     aux_loss=True, targets=None, with stand-in backbones that hand the token embeddings through; name-seeded
     parameters; pred_logits / pred_boxes / aux outputs / hs_det
"""
import importlib.util
import os

import numpy as np
import torch

from make_golden_transformer import randomise, randomise_by_name

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src/trackformer/models/transformer.py"


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_kine_transformer", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def tokens(g, N, T, Lq, d, n_track, pad):
    src_b = torch.randn(N, T, d, generator=g).requires_grad_(True)
    src_m = torch.randn(N, T, d, generator=g).requires_grad_(True)
    pos_b = 0.3 * torch.randn(N, T, d, generator=g)
    pos_m = 0.3 * torch.randn(N, T, d, generator=g)
    mask = torch.zeros(N, T, dtype=torch.bool)
    if pad:
        mask[N - 1, T - pad:] = True
    qe_b = (0.5 * torch.randn(Lq, N, d, generator=g)).requires_grad_(True)
    qe_m = (0.5 * torch.randn(Lq, N, d, generator=g)).requires_grad_(True)
    tgt_b = torch.zeros(Lq, N, d)
    tgt_m = torch.zeros(Lq, N, d)
    tgt_b[:n_track] = torch.randn(n_track, N, d, generator=g)
    tgt_m[:n_track] = torch.randn(n_track, N, d, generator=g)
    return src_b, src_m, mask, qe_b, qe_m, tgt_b, tgt_m, pos_b, pos_m


def main():
    tr = load_reference()
    blob = {}
    g = torch.Generator().manual_seed(33)

    # ---- case a ----
    d, N, T, Lq = 72, 2, 11, 7
    model = tr.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=2, num_decoder_layers=2,
                                        dim_feedforward=96, dropout=0.0, return_intermediate_dec=True)
    randomise(model, g)
    model.eval()
    src_b, src_m, mask, qe_b, qe_m, tgt_b, tgt_m, pos_b, pos_m = tokens(g, N, T, Lq, d, 2, 3)
    hs_det, hs_meta, raw_det, mem_det = model(src_b, src_m, mask, qe_b, qe_m, tgt_b, tgt_m, pos_b, pos_m)
    w1 = torch.randn(hs_det.shape, generator=g)
    w2 = torch.randn(hs_meta.shape, generator=g)
    w3 = torch.randn(mem_det.shape, generator=g)
    loss = (hs_det * w1).sum() + (hs_meta * w2).sum() + (mem_det * w3).sum() + (raw_det * w1).sum() * 0.5
    params = dict(model.named_parameters())
    names = [k for k, p in params.items() if "linear2" not in k or "branch" not in k]   # IntertwinedBranch.linear2 is unused
    grads = torch.autograd.grad(loss, [src_b, src_m, qe_b, qe_m] + [params[k] for k in names])
    blob.update({"a:src_b": src_b, "a:src_m": src_m, "a:mask": mask, "a:qe_b": qe_b, "a:qe_m": qe_m, "a:tgt_b": tgt_b,
                 "a:tgt_m": tgt_m, "a:pos_b": pos_b, "a:pos_m": pos_m, "a:w1": w1, "a:w2": w2, "a:w3": w3,
                 "a:hs_det": hs_det, "a:hs_meta": hs_meta, "a:raw_det": raw_det, "a:mem_det": mem_det,
                 "a:g_src_b": grads[0], "a:g_src_m": grads[1], "a:g_qe_b": grads[2], "a:g_qe_m": grads[3]})
    for k, p in params.items():
        blob["a:p:" + k] = p
    for k, gk in zip(names, grads[4:]):
        if gk.dim() == 1 or ".layers.0." in k or "branch" in k:
            blob["a:g:" + k] = gk

    # ---- case b ----
    d, N, T, Lq = 288, 1, 23, 160
    model = tr.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                        dim_feedforward=64, dropout=0.0, return_intermediate_dec=True)
    randomise_by_name(model)
    model.eval()
    args = tokens(g, N, T, Lq, d, 10, 0)
    with torch.no_grad():
        out = model(*args)
    for k, v in zip(("src_b", "src_m", "mask", "qe_b", "qe_m", "tgt_b", "tgt_m", "pos_b", "pos_m"), args):
        blob["b:" + k] = v
    for k, v in zip(("hs_det", "hs_meta", "raw_det", "mem_det"), out):
        blob["b:" + k] = v

    # ---- case c ----
    d, N, T = 72, 2, 13
    model = tr.DualKinematicEncoder(d_model=d, nhead=8, num_encoder_layers=2, dim_feedforward=48, dropout=0.0)
    randomise_by_name(model)
    model.eval()
    src_b, src_m, mask, _, _, _, _, pos_b, pos_m = tokens(g, N, T, 1, d, 0, 4)
    with torch.no_grad():
        out = model(src_b, src_m, mask, pos_b, pos_m)
    for k, v in (("src_b", src_b), ("src_m", src_m), ("mask", mask), ("pos_b", pos_b), ("pos_m", pos_m)):
        blob["c:" + k] = v
    for k, v in zip(("hs_det", "hs_meta", "mem_meta", "mem_det"), out):
        blob["c:" + k] = v

    # ---- case d ----
    d, N, T, Lq = 72, 2, 9, 5
    model = tr.KinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2, dim_feedforward=48,
                                    dropout=0.0, normalize_before=True, return_intermediate_dec=False)
    randomise_by_name(model)
    model.eval()
    src_b, _, mask, qe_b, _, tgt_b, _, pos_b, _ = tokens(g, N, T, Lq, d, 1, 2)
    with torch.no_grad():
        out = model(src_b, mask, qe_b, tgt_b, pos_b)
    for k, v in (("src", src_b), ("mask", mask), ("qe", qe_b), ("tgt", tgt_b), ("pos", pos_b)):
        blob["d:" + k] = v
    for k, v in zip(("hs", "raw", "memory"), out):
        blob["d:" + k] = v

    # ---- case e ----
    from make_golden_layers import load_reference as load_pkg
    load_pkg()
    import importlib
    detr = importlib.import_module("trackformer.models.detr")

    class Feat:
        def __init__(self, x, m):
            self.x, self.m = x, m

        def decompose(self):
            return self.x, self.m

    class Samples:
        pass
    d, N, T, Lq = 72, 2, 8, 6
    src_b, src_m, mask, _, _, _, _, pos_b, pos_m = tokens(g, N, T, 1, d, 0, 2)
    back = [lambda x: ([Feat(x[0], mask)], [x[1]])] * 2
    transformer = tr.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    model = detr.KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=Lq, aux_loss=True,
                                              dim_tracklets_det=16, dim_tracklets_metadata=8)
    randomise_by_name(model)
    model.eval()
    smp = Samples()
    smp.detections, smp.metadata = (src_b.detach(), pos_b), (src_m.detach(), pos_m)
    with torch.no_grad():
        out, _, _, src_out, hs_det = model(smp, None)
    for k, v in (("src_b", src_b), ("src_m", src_m), ("mask", mask), ("pos_b", pos_b), ("pos_m", pos_m),
                 ("pred_logits", out["pred_logits"]), ("pred_boxes", out["pred_boxes"]),
                 ("aux_logits", out["aux_outputs"][0]["pred_logits"]), ("aux_boxes", out["aux_outputs"][0]["pred_boxes"]),
                 ("hs_det", hs_det), ("src_out", src_out)):
        blob["e:" + k] = v

    path = os.path.join(HERE, "kine_f32.npz")
    np.savez_compressed(path, **{k: v.detach().numpy() for k, v in blob.items()})
    print("wrote", path, os.path.getsize(path), "bytes;", len(blob), "arrays")


if __name__ == "__main__":
    main()
