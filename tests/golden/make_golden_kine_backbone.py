#!/usr/bin/env python
"""Generate tests/golden/kine_backbone_f32.npz from the REAL reference `kine` backbones and detector.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_kine_backbone.py

Imports (package __init__ files bypassed, visdom stubbed: make_golden_layers.load_reference) the reference's
src/trackformer/models/backbone.py (Kinet_Backbone, LayerBackboneFC, Joiner :110-187), position_encoding.py
(PositionEmbeddingSineDetections :152-185), util/misc.py (NestedTensor), detr.py (KinematicDetectorTransformer) and
transformer.py (DualKinematicTransformer), fp32 on the CPU, eval mode; parameters are a function of their names
(make_golden_transformer.randomise_by_name), not stored.

  f: the two Joiners as build_backbone wires them (backbone.py:188-212) for hidden_dim 72: detections (BS 2, T 9, 4 box
     coordinates; 3 of the second sample's detections padded) and metadata (confidence); max_detections 7, so the
     positional rank wraps around.  Stored: inputs, mask, token embeddings and positional embeddings of both.
  g: KinematicDetectorTransformer over those Joiners and a d_model 72, 1 + 2 layer DualKinematicTransformer, 6 queries,
     aux_loss=True, targets=None -- the whole `kine` model from raw detections: pred_logits / pred_boxes / hs_det.
"""
import importlib
import os

import numpy as np
import torch

from make_golden_layers import load_reference as load_pkg
from make_golden_transformer import randomise_by_name

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    load_pkg()
    bb = importlib.import_module("trackformer.models.backbone")
    pe = importlib.import_module("trackformer.models.position_encoding")
    misc = importlib.import_module("trackformer.util.misc")
    detr = importlib.import_module("trackformer.models.detr")
    tr = importlib.import_module("trackformer.models.transformer")
    g = torch.Generator().manual_seed(57)
    d, N, T = 72, 2, 9
    boxes = torch.rand(N, T, 4, generator=g)
    conf = torch.rand(N, T, 1, generator=g)
    mask = torch.zeros(N, T, dtype=torch.bool)
    mask[1, T - 3:] = True
    det = bb.Kinet_Backbone(4, hidden_dims=[256, 512, d], activation="relu", return_interm_layers=False)
    meta = bb.Kinet_Backbone(1, hidden_dims=[16, 64, d], activation="relu", return_interm_layers=False)
    emb = pe.PositionEmbeddingSineDetections(d // 2, max_detections=7)
    back = [bb.Joiner(det, emb), bb.Joiner(meta, emb)]
    for b in back:
        randomise_by_name(b)
        b.eval()
    blob = {"f:boxes": boxes, "f:conf": conf, "f:mask": mask}
    with torch.no_grad():
        for name, b, x in (("det", back[0], boxes), ("meta", back[1], conf)):
            feats, pos = b(misc.NestedTensor(x, mask))
            assert len(feats) == 1
            blob["f:tokens_" + name] = feats[0].tensors
            blob["f:pos_" + name] = pos[0]

    transformer = tr.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    model = detr.KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=6, aux_loss=True,
                                              dim_tracklets_det=16, dim_tracklets_metadata=8)
    randomise_by_name(model)
    model.eval()
    smp = misc.NestedTensorKinet(misc.NestedTensor(boxes, mask), misc.NestedTensor(conf, mask))
    with torch.no_grad():
        out, _, _, src_out, hs_det = model(smp, None)
    for k, v in (("pred_logits", out["pred_logits"]), ("pred_boxes", out["pred_boxes"]),
                 ("aux_logits", out["aux_outputs"][0]["pred_logits"]), ("hs_det", hs_det), ("src_out", src_out)):
        blob["g:" + k] = v
    path = os.path.join(HERE, "kine_backbone_f32.npz")
    np.savez_compressed(path, **{k: v.detach().numpy() for k, v in blob.items()})
    print("wrote", path, os.path.getsize(path), "bytes;", len(blob), "arrays")


if __name__ == "__main__":
    main()
