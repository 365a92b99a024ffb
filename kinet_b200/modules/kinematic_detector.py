"""KinematicDetectorTransformer -- the KineT detector head around DualKinematicTransformer (SURVEY.md section 8 row
f4), mirroring the reference's src/trackformer/models/detr.py:288-417: same constructor arguments, sub-module /
parameter names (``class_embed``, ``bbox_embed``, ``query_embed_det``, ``query_embed_metadata``,
``input_proj_tracklets_det``, ``input_proj_tracklets_metadata``, ``backbone_det``, ``backbone_metadata``), forward
signature, ``targets`` keys and return tuple.

The two backbones are the caller's (the reference's ``Joiner(Kinet_Backbone, position_embedding)`` pair,
backbone.py:128-215, out of scope here): anything callable as ``features, pos = backbone(tensors)`` whose
``features[-1].decompose()`` yields ``(tokens (BS, T, C), mask (BS, T))`` and ``pos[0]`` the positional embedding.
``TokenFeatures`` below is the minimal such pair for harnesses that already hold token embeddings.

What runs differently from the reference: the transformer is kinet_b200's (fused add + LayerNorm kernels, the
short-sequence attention kernel); the heads' Linear + ReLU stacks go through the GEMM-epilogue ReLU of
functions/layer_ops.py; the per-sample ``torch.stack`` / ``zeros`` / ``cat`` glue of :366-395 is kept as is (it is
tiny and only runs when track queries are handed in through ``targets``).
"""
import torch
from torch import nn

from ..functions.layer_ops import linear


class MLP(nn.Module):
    """The reference's MLP (detr.py:937-951): Linear-ReLU stack, ReLU in the GEMM epilogue."""

    def __init__(self, input_dim, hidden_dim, output_dim, num_layers):
        super().__init__()
        self.num_layers = num_layers
        h = [hidden_dim] * (num_layers - 1)
        self.layers = nn.ModuleList(nn.Linear(n, k) for n, k in zip([input_dim] + h, h + [output_dim]))

    def forward(self, x):
        for i, layer in enumerate(self.layers):
            x = linear(x, layer, relu=i < self.num_layers - 1)
        return x


class TokenFeatures:
    """A (tokens, mask) pair with the reference's NestedTensor.decompose() interface."""

    def __init__(self, tensors, mask):
        self.tensors, self.mask = tensors, mask

    def decompose(self):
        return self.tensors, self.mask


class KinematicDetectorTransformer(nn.Module):
    def __init__(self, backbone, transformer, num_classes, num_queries, aux_loss=False, overflow_boxes=False,
                 dim_tracklets_det=128, dim_tracklets_metadata=8):
        super().__init__()
        self.num_queries = num_queries
        self.transformer = transformer
        self.overflow_boxes = overflow_boxes
        self.class_embed = nn.Linear(self.hidden_dim, num_classes + 1)
        self.bbox_embed = MLP(self.hidden_dim, self.hidden_dim, 4, 3)
        self.query_embed_det = nn.Embedding(num_queries, self.hidden_dim)
        self.query_embed_metadata = nn.Embedding(num_queries, self.hidden_dim)
        self.input_proj_tracklets_det = MLP(dim_tracklets_det, self.hidden_dim, self.hidden_dim, 3)
        self.input_proj_tracklets_metadata = MLP(dim_tracklets_metadata, self.hidden_dim // 2, self.hidden_dim, 3)
        self.backbone_det = backbone[0]
        self.backbone_metadata = backbone[1]
        self.aux_loss = aux_loss

    @property
    def hidden_dim(self):
        return self.transformer.d_model

    def forward(self, samples, targets=None):
        """samples.detections / samples.metadata feed the two backbones (:349-350).  Returns the reference's tuple
        ``(out, targets, features_det, src_det, hs_det)`` with out['pred_logits'] (BS, Lq, num_classes + 1) and
        out['pred_boxes'] (BS, Lq, 4) (+ 'aux_outputs' with aux_loss)."""
        features_det, pos_det = self.backbone_det(samples.detections)
        features_metadata, pos_metadata = self.backbone_metadata(samples.metadata)
        src_det, mask = features_det[-1].decompose()
        src_metadata, _ = features_metadata[-1].decompose()
        batch_size = src_det.size()[0]

        query_embed_det = self.query_embed_det.weight.unsqueeze(1).repeat(1, batch_size, 1)
        query_embed_metadata = self.query_embed_metadata.weight.unsqueeze(1).repeat(1, batch_size, 1)
        tgt_det = tgt_metadata = None
        if targets is not None and len(targets[0]['track_query_hs_embeds_det']) > 0:        # :364-395
            hs_det_in = torch.stack([t['track_query_hs_embeds_det'] for t in targets])
            n_track = hs_det_in.shape[1]
            zeros = torch.zeros(n_track, batch_size, self.hidden_dim, device=query_embed_det.device,
                                dtype=query_embed_det.dtype)
            query_embed_det = torch.cat([zeros, query_embed_det], dim=0)
            tgt_det = torch.zeros_like(query_embed_det)
            tgt_det[:n_track] = self.input_proj_tracklets_det(hs_det_in.transpose(0, 1)).to(tgt_det.dtype)
            hs_meta_in = torch.stack([t['track_query_hs_embeds_meta'] for t in targets])
            # (the reference concatenates the already track-padded query_embed_det here, :389-391, which gives the
            # metadata branch n_track more queries than the detection branch and fails in IntertwinedBranch's
            # concatenation: one of the work-in-progress paths of SURVEY.md F7.  The evident intent is built instead.)
            query_embed_metadata = torch.cat([zeros, query_embed_metadata], dim=0)
            tgt_metadata = torch.zeros_like(query_embed_metadata)
            tgt_metadata[:n_track] = self.input_proj_tracklets_metadata(hs_meta_in.transpose(0, 1)).to(tgt_metadata.dtype)
        hs_det, hs_metadata, hs_without_norm_det, memory_det = self.transformer(
            src_det, src_metadata, mask, query_embed_det, query_embed_metadata, tgt_det, tgt_metadata,
            pos_boxes=pos_det[0], pos_metadata=pos_metadata[0])

        outputs_class = linear(hs_metadata, self.class_embed)
        outputs_coord = self.bbox_embed(hs_det).sigmoid()
        out = {'pred_logits': outputs_class[-1], 'pred_boxes': outputs_coord[-1]}
        if self.aux_loss:
            out['aux_outputs'] = self._set_aux_loss(outputs_class, outputs_coord)
        return out, targets, features_det, src_det, hs_det

    @torch.jit.unused
    def _set_aux_loss(self, outputs_class, outputs_coord):
        return [{'pred_logits': a, 'pred_boxes': b} for a, b in zip(outputs_class[:-1], outputs_coord[:-1])]
