"""Per-frame tracking loop around kinet_b200's deformable transformer -- the caller the headline metric
("KineT frames/sec") is named after, SURVEY.md section 8(f) row 3 and BASELINE.json configs[2].

Mirrors the reference's per-frame logic
  * src/trackformer/models/tracker.py:269-309  -- the live tracks of frame t become frame t+1's track queries
    (content = the track's last decoder output ``hs_embed``, reference point = its box centre),
  * :334-420 -- a track survives while its score stays above ``track_obj_score_thresh`` (terminated after
    ``steps_termination`` consecutive misses), NMS among tracks,
  * :422-505 -- object queries above ``detection_obj_score_thresh`` start new tracks; NMS in which existing tracks
    always win over new ones,
  * :536-557 -- per-frame results, and src/track.py:127-161 -- the frame loop over a sequence,
  * src/trackformer/models/deformable_detr.py:229-252, 286-326 -- class / box heads on the last decoder layer
    (box = sigmoid(MLP(hs) + inverse_sigmoid(reference))), scores = sigmoid(logits).max(-1), boxes scaled to pixels,
on top of ``kinet_b200.modules.DeformableTransformer`` -- eagerly, or through ``FrameGraphs`` (one CUDA graph for
the encoder, one per bucket of query counts for the decoder): the number of queries changes EVERY frame.

What is different from the reference, on purpose: track state lives in device tensors (boxes, embeddings) instead
of one Python object per track holding device scalars, and a frame is read back ONCE (scores, labels, boxes of
all queries in one packed copy) instead of once per track and field (tracker.py:293, 536-542); thresholding and
NMS then run on the host on a few hundred boxes.  Re-identification through inactive track queries (:273-276,
:290-304, :367-394; ``inactive_patience`` >= 0 -- cfgs/track.yaml sets -1, i.e. off) is built, and so is ``reid()`` (:170-267: greedy centre-distance or embedding-distance Hungarian matching of new
detections to inactive tracks); not covered: masks, attention maps -- they contain no MSDeformAttn work.  The backbone is out of scope: frames
enter as the multi-scale feature maps the backbone + input projections would produce.
"""
import math

from collections.abc import Mapping

import numpy as np
import torch
from torch import nn

from .frame_graph import FrameGraphs
from .modules.deformable_transformer import _shapes_tensor, inverse_sigmoid


class MLP(nn.Module):
    """The reference's box head (src/trackformer/models/detr.py MLP: Linear-ReLU stack)."""

    def __init__(self, input_dim, hidden_dim, output_dim, num_layers):
        super().__init__()
        h = [hidden_dim] * (num_layers - 1)
        self.layers = nn.ModuleList(nn.Linear(n, k) for n, k in zip([input_dim] + h, h + [output_dim]))

    def forward(self, x):
        for i, layer in enumerate(self.layers):
            x = torch.relu(layer(x)) if i < len(self.layers) - 1 else layer(x)
        return x


class DetectionHeads(nn.Module):
    """class_embed / bbox_embed of the last decoder layer (deformable_detr.py:88-110, 229-252; inference has no use
    for the auxiliary heads)."""

    def __init__(self, d_model, num_classes=1):
        super().__init__()
        self.class_embed = nn.Linear(d_model, num_classes)
        self.bbox_embed = MLP(d_model, d_model, 4, 3)
        prior = 0.01
        nn.init.constant_(self.class_embed.bias, -math.log((1 - prior) / prior))
        nn.init.constant_(self.bbox_embed.layers[-1].weight, 0.0)
        nn.init.constant_(self.bbox_embed.layers[-1].bias, 0.0)

    def forward(self, hs_last, reference):
        """hs_last (N, Lq, C); reference (N, Lq, 2|4) in [0,1] -> logits (N, Lq, K), boxes (N, Lq, 4) cxcywh in [0,1]."""
        logits = self.class_embed(hs_last.float())
        tmp = self.bbox_embed(hs_last.float())
        ref = inverse_sigmoid(reference.float())
        if ref.shape[-1] == 4:
            tmp = tmp + ref
        else:
            tmp = torch.cat([tmp[..., :2] + ref, tmp[..., 2:]], -1)
        return logits, tmp.sigmoid()


def _iou_matrix(boxes):
    """(n, n) IoU of xyxy boxes in float32 (what torchvision.ops.nms computes in), from contiguous coordinate rows."""
    x1, y1, x2, y2 = np.ascontiguousarray(boxes.T, dtype=np.float32)
    area = np.maximum(x2 - x1, 0) * np.maximum(y2 - y1, 0)
    iw = np.minimum(x2[:, None], x2) - np.maximum(x1[:, None], x1)
    ih = np.minimum(y2[:, None], y2) - np.maximum(y1[:, None], y1)
    inter = np.maximum(iw, 0, out=iw)
    inter *= np.maximum(ih, 0, out=ih)
    union = area[:, None] + area
    union -= inter
    return np.divide(inter, np.maximum(union, 1e-12, out=union), out=inter)


def _greedy_keep(over, order):
    """Greedy suppression over a boolean (n, n) 'i suppresses j' matrix (diagonal False) visiting ``order``: the sorted
    indices kept."""
    dead = np.zeros(len(over), dtype=bool)
    keep = []
    for i in order:
        if not dead[i]:
            keep.append(i)
            dead |= over[i]
    return np.asarray(sorted(keep), dtype=np.int64)


def _nms(boxes, scores, thresh):
    """Indices kept by greedy IoU suppression (torchvision.ops.nms semantics: IoU > thresh suppresses), on the host.
    One vectorised IoU matrix; the common case (nothing overlaps that much) returns without a Python loop."""
    n = len(boxes)
    if n <= 1:
        return np.arange(n, dtype=np.int64)
    over = _iou_matrix(boxes) > thresh
    over.flat[::n + 1] = False
    if not over.any():
        return np.arange(n, dtype=np.int64)
    return _greedy_keep(over, np.argsort(-scores, kind="stable"))


def public_detections_mask(new_boxes, public_boxes, mode):
    """The reference's Tracker.public_detections_mask (tracker.py:126-167) on host arrays: which of the frame's new
    detections (n, 4) xyxy are backed by one of the sequence's PUBLIC detections (m, 4) xyxy.  ``mode``: False / None = all;
    'center_distance': per public box the nearest unclaimed detection, if the squared centre distance is below that
    detection's area; 'min_iou_0_5': per public box the unclaimed detection with the largest IoU, if >= 0.5."""
    n = len(new_boxes)
    if not mode:
        return np.ones(n, dtype=bool)
    mask = np.zeros(n, dtype=bool)
    if public_boxes is None or not len(public_boxes) or not n:
        return mask
    a, b = np.asarray(new_boxes, np.float64), np.asarray(public_boxes, np.float64)
    if mode == "center_distance":
        size = ((a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])).astype(np.float32)
        ca, cb = (a[:, :2] + a[:, 2:]) / 2, (b[:, :2] + b[:, 2:]) / 2
        dist = ((ca[:, None] - cb[None]) ** 2).sum(2)
        for j in range(len(b)):
            i = int(dist[:, j].argmin())
            if dist[i, j] < size[i]:
                dist[i, :] = 1e18
                mask[i] = True
    elif mode == "min_iou_0_5":
        lt, rb = np.maximum(a[:, None, :2], b[None, :, :2]), np.minimum(a[:, None, 2:], b[None, :, 2:])
        inter = np.clip(rb - lt, 0, None).prod(2)
        area_a, area_b = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1]), (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
        iou = inter / np.maximum(area_a[:, None] + area_b[None] - inter, 1e-12)
        for j in range(len(b)):
            i = int(iou[:, j].argmax())
            if iou[i, j] >= 0.5:
                iou[i, :] = 0
                mask[i] = True
    else:
        raise NotImplementedError(mode)
    return mask


class SequenceTracker:
    """Tracks one sequence frame by frame.

        tracker = SequenceTracker(transformer, heads, query_embed, img_size=(1080, 1920))
        for srcs, masks, pos_embeds in frames:           # per level (1, C, H, W) / (1, H, W) bool / (1, C, H, W)
            result = tracker.step(srcs, masks, pos_embeds)   # {track_id: {"bbox": xyxy px, "score": float}}

    transformer: kinet_b200.modules.DeformableTransformer (not two-stage); heads: DetectionHeads;
    query_embed: (num_object_queries, 2C) learned object queries.  use_graphs: replay FrameGraphs (the query count
    is padded to ``bucket``); otherwise the transformer runs eagerly.  autocast_dtype: None = fp32."""

    def __init__(self, transformer, heads, query_embed, img_size, detection_obj_score_thresh=0.4,
                 track_obj_score_thresh=0.4, detection_nms_thresh=0.9, track_nms_thresh=0.9, steps_termination=1,
                 use_graphs=False, bucket=32, autocast_dtype=None, inactive_patience=-1, reid_score_thresh=0.4,
                 public_detections=False, reid_sim_threshold=0.0, reid_greedy_matching=False):
        assert not transformer.two_stage and not transformer.multi_frame_attention_separate_encoder
        # reid() (tracker.py:170-267): new detections are first offered to the inactive tracks -- greedily by centre distance
        # (reid_greedy_matching) or by Hungarian assignment on the distance of the decoder embeddings, accepted below
        # reid_sim_threshold.  (At the configured threshold 0.0 the embedding variant never accepts; it is skipped then.)
        self.reid_sim_threshold, self.reid_greedy = reid_sim_threshold, reid_greedy_matching
        # public_detections: False | 'center_distance' | 'min_iou_0_5' (tracker.py:126-167): new tracks only where one of the
        # frame's public detections (step(..., public_dets=(m, 4) xyxy pixels)) backs the detection
        self.public_detections = public_detections
        # inactive_patience >= 0: a terminated track stays a track query (its last box and embedding) for that many frames
        # and returns under its id when that query scores above reid_score_thresh (tracker.py:273-276, :290-304, :367-383);
        # -1 (cfgs/track.yaml:44) forgets it at once.  The embedding-distance reid() (:170-267) is not built: at the
        # configured reid_sim_threshold 0.0 it never matches.
        self.inactive_patience, self.reid_thresh = inactive_patience, reid_score_thresh
        self.transformer, self.heads = transformer, heads
        self.query_embed = query_embed
        self.num_object_queries = query_embed.shape[0]
        self.img_h, self.img_w = float(img_size[0]), float(img_size[1])
        self.det_thresh, self.trk_thresh = detection_obj_score_thresh, track_obj_score_thresh
        self.det_nms, self.trk_nms = detection_nms_thresh, track_nms_thresh
        self.steps_termination = steps_termination
        self.use_graphs, self.bucket, self.autocast_dtype = use_graphs, bucket, autocast_dtype
        self._graphs = None
        self.reset()

    def reset(self):
        """Start a new sequence (tracker.py:75-95); recorded graphs are kept (same weights, same pyramid)."""
        self.track_ids = []                      # host: ids of the live tracks, in query order
        self.track_misses = []                   # host: consecutive frames below track_obj_score_thresh
        self.inactive_ids = []                   # host: ids of the inactive tracks, in query order behind the live ones
        self.inactive_count = []                 # host: frames spent inactive
        self.track_boxes = None                  # device (T + I, 4) xyxy pixels: live tracks, then inactive ones
        self.track_scores = None                 # host numpy (T + I,)
        self._host_boxes = None                  # host numpy (T + I, 4), the same boxes
        self.track_hs = None                     # device (T + I, C)
        self.num_reids = 0
        self.next_id = 0
        self.frame_index = 0
        self.results = {}

    # ---- the transformer call ---------------------------------------------------------------------------------------
    def _flatten(self, srcs, masks, pos_embeds):
        tr = self.transformer
        shapes = [(int(src.shape[2]), int(src.shape[3])) for src in srcs]
        src = torch.cat([s_.flatten(2).transpose(1, 2) for s_ in srcs], 1)
        # the positional maps (+ level embedding) are the same tensors frame after frame: flattened once
        key = tuple((id(p), p._version, p.data_ptr()) for p in pos_embeds) + (tr.level_embed._version,)
        hit = getattr(self, "_pos_flat", None)
        if hit is None or hit[0] != key:
            pos = torch.cat([p.flatten(2).transpose(1, 2) + tr.level_embed[lvl].view(1, 1, -1)
                             for lvl, p in enumerate(pos_embeds)], 1)
            hit = self._pos_flat = (key, pos, list(pos_embeds))      # (the list keeps the ids alive)
        return src, hit[1], shapes

    def _run_transformer(self, srcs, masks, pos_embeds, targets):
        tr = self.transformer
        if not self.use_graphs:
            with torch.autocast("cuda", dtype=self.autocast_dtype, enabled=self.autocast_dtype is not None):
                hs, _, init_ref, inter_ref, _, _ = tr(srcs, masks, pos_embeds, self.query_embed, targets)
            if hs.dim() == 4:                    # return_intermediate_dec: last layer, reference of the layer before
                ref = inter_ref[-2] if inter_ref.shape[0] > 1 else init_ref
                return hs[-1], ref
            return hs, init_ref
        # graphs: the pyramid is fixed, the number of queries is not.  No padding in tracking frames (N = 1).
        self._encode_graph(srcs, masks, pos_embeds)
        return self._decode_graph(targets)

    def _encode_graph(self, srcs, masks, pos_embeds):
        tr = self.transformer
        src, pos, shapes = self._flatten(srcs, masks, pos_embeds)
        dev = src.device
        if self._graphs is None:
            self._graphs = FrameGraphs(tr.encoder, tr.decoder, _shapes_tensor(shapes, dev), bucket=self.bucket,
                                       autocast_dtype=self.autocast_dtype, post=self._graph_post)
            self._valid = torch.ones(1, len(shapes), 2, device=dev)
        self._graphs.encode(src, pos, self._valid)

    def _decode_graph(self, targets):
        tr = self.transformer
        c = self.query_embed.shape[1] // 2
        qpos, tgt = torch.split(self.query_embed, c, dim=1)
        qpos, tgt = qpos[None], tgt[None]
        ref = tr.reference_points(qpos).sigmoid()
        if targets is not None:
            prev_hs = targets[0]["track_query_hs_embeds"][None]
            qpos = torch.cat([torch.zeros_like(prev_hs), qpos], 1)
            tgt = torch.cat([prev_hs, tgt], 1)
            ref = torch.cat([targets[0]["track_query_boxes"][None, :, :2], ref], 1)
        hs, ref_out, _ = self._graphs.decode(tgt, ref, qpos)
        post = self._graphs.post_out
        if post is not None:             # heads + packing were recorded behind the decoder: slice the padded rows off
            Lq = hs.shape[-2]
            self._pre = (post[1][:, :Lq], post[2][:Lq])
            return post[0][:, :Lq], None
        if hs.dim() == 4:
            return hs[-1], (ref_out[-2] if ref_out.shape[0] > 1 else ref)
        return hs, ref

    def _graph_post(self, hs, ref_out):
        """Recorded into the decoder graphs (FrameGraphs post): class / box heads and the packed read-back rows of the
        last decoder layer -- ~20 small kernels that would otherwise be launched one by one from Python every frame."""
        if hs.dim() != 4 or ref_out.shape[0] < 2:
            return None                  # the reference of the last layer is the caller's: heads run eagerly
        return self._heads_and_pack(hs[-1], ref_out[-2])

    # ---- one frame --------------------------------------------------------------------------------------------------
    def _targets(self, dev):
        if getattr(self, "_scale", None) is None or self._scale.device != dev:
            self._scale = torch.tensor([self.img_w, self.img_h, self.img_w, self.img_h], device=dev)
        self._prune_inactive(dev)
        if not self.track_ids and not self.inactive_ids:
            return None
        b = self.track_boxes                                       # xyxy px -> cxcywh normalised (tracker.py:299-304)
        cxcywh = torch.stack([(b[:, 0] + b[:, 2]) / 2, (b[:, 1] + b[:, 3]) / 2, b[:, 2] - b[:, 0], b[:, 3] - b[:, 1]], -1)
        return [{"track_query_boxes": cxcywh / self._scale, "track_query_hs_embeds": self.track_hs}]

    def _prune_inactive(self, dev):
        """Start of a frame (tracker.py:273-276): inactive tracks beyond their patience, or without area, are forgotten."""
        I = len(self.inactive_ids)
        if not I:
            return
        T = len(self.track_ids)
        hb = self._host_boxes[T:]
        ok = [self.inactive_count[j] <= self.inactive_patience and hb[j, 2] > hb[j, 0] and hb[j, 3] > hb[j, 1] for j in range(I)]
        if all(ok):
            return
        rows = list(range(T)) + [T + j for j in range(I) if ok[j]]
        self.inactive_ids = [t for t, k in zip(self.inactive_ids, ok) if k]
        self.inactive_count = [c for c, k in zip(self.inactive_count, ok) if k]
        self._host_boxes, self.track_scores = self._host_boxes[rows], self.track_scores[rows]
        if rows:
            (idx,) = self._to_device_indices(dev, rows)
            self.track_boxes, self.track_hs = self.track_boxes[idx], self.track_hs[idx]
        else:
            self.track_boxes = self.track_hs = None

    @torch.no_grad()
    def step(self, srcs, masks, pos_embeds, public_dets=None):
        """One frame, start to finish (the device idles while the host reads the results and updates the tracks;
        ``run`` overlaps that with the next frame's encoder)."""
        dev = srcs[0].device
        hs, ref = self._run_transformer(srcs, masks, pos_embeds, self._targets(dev))
        hs, xyxy, packed = self._heads_and_pack(hs, ref)
        return self._update_tracks(hs, xyxy, packed.cpu().numpy(), public_dets)   # ONE device -> host read per frame

    @torch.no_grad()
    def run(self, frames):
        """(frames: (srcs, masks, pos_embeds) or, with public detections, (srcs, masks, pos_embeds, public_dets).)
        Generator over a sequence, ``for result in tracker.run(frames)``: like calling step() on every frame, but
        with the CUDA-graph path frame t+1's encoder (which does not depend on the tracks) is enqueued BEFORE the host
        waits for frame t's results, so the read-back, the thresholding / NMS and the launches of the next decoder
        hide behind ~60 % of a frame's device work.  Results are identical to step()'s."""
        if not self.use_graphs:
            for frame in frames:
                yield self.step(*frame)
            return
        it = iter(frames)
        try:
            cur = next(it)
        except StopIteration:
            return
        dev = cur[0][0].device
        side = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)
        host_buf = None
        self._encode_graph(*cur[:3])
        while cur is not None:
            hs, ref = self._decode_graph(self._targets(dev))
            hs, xyxy, packed = self._heads_and_pack(hs, ref)
            if host_buf is None or host_buf.shape[0] < packed.shape[0]:
                host_buf = torch.empty(packed.shape[0] + 256, 6, dtype=packed.dtype).pin_memory()
            ready = torch.cuda.Event()
            ready.record(main)
            side.wait_event(ready)
            with torch.cuda.stream(side):
                host_buf[:packed.shape[0]].copy_(packed, non_blocking=True)
                copied = torch.cuda.Event()
                copied.record(side)
            packed.record_stream(side)
            nxt = next(it, None)
            if nxt is not None:
                self._encode_graph(*nxt[:3])      # the device keeps working while the host reads frame t back
            copied.synchronize()
            yield self._update_tracks(hs, xyxy, host_buf[:packed.shape[0]].numpy().copy(), cur[3] if len(cur) > 3 else None)
            cur = nxt

    def _heads_and_pack(self, hs, ref):
        pre = getattr(self, "_pre", None)
        if ref is None and pre is not None:      # done inside the decoder graph (_graph_post)
            self._pre = None
            return hs, pre[0], pre[1]
        logits, boxes = self.heads(hs, ref)
        prob = logits.sigmoid()
        scores, labels = prob.max(-1)
        x, y, w, h = boxes.unbind(-1)
        xyxy = torch.stack([x - 0.5 * w, y - 0.5 * h, x + 0.5 * w, y + 0.5 * h], -1)
        xyxy = torch.minimum((xyxy * self._scale).clamp(min=0), self._scale)     # scale to pixels, clip_boxes_to_image
        packed = torch.cat([scores[0, :, None], labels[0, :, None].to(scores.dtype), xyxy[0]], -1)
        return hs, xyxy, packed

    def _to_device_indices(self, dev, *lists):
        """Index lists -> device int64 tensors through PINNED staging memory and one non-blocking copy.  A copy from
        pageable memory (torch.as_tensor(list, device=...)) first synchronises the stream it is issued on -- here the
        stream that already holds the next frame's encoder, i.e. the host would sit out the very work run() queued to
        hide its own latency.  Two staging buffers alternate: a frame's copy has executed (it is ordered before that
        frame's decoder, whose results the host has read) long before its buffer comes round again."""
        n = sum(len(l) for l in lists)
        st = getattr(self, "_idx_stage", None)
        if st is None or st[0].shape[0] < n or st[2].device != dev:
            cap = max(1024, 2 * n)
            st = [torch.empty(cap, dtype=torch.int64).pin_memory(), torch.empty(cap, dtype=torch.int64).pin_memory(),
                  torch.empty(cap, dtype=torch.int64, device=dev), 0]
            self._idx_stage = st
        st[3] ^= 1
        host = st[st[3]]
        flat = host.numpy()
        out, o = [], 0
        for l in lists:
            flat[o:o + len(l)] = l
            out.append((o, o + len(l)))
            o += len(l)
        st[2][:n].copy_(host[:n], non_blocking=True)
        return [st[2][a:b] for a, b in out]      # (consumed by this frame's gathers, stream-ordered before the next copy)

    def _reid(self, hs, live, inactive, new):
        """tracker.py:170-267 on the entry lists of _update_tracks: matched inactive tracks return (this frame's detection
        is their new state, appended behind the live tracks), their detections start no new track."""
        ib = np.stack([e[4] for e in inactive])
        nb = np.stack([e[4] for e in new])
        if self.reid_greedy:
            ci, cn = (ib[:, :2] + ib[:, 2:]) / 2, (nb[:, :2] + nb[:, 2:]) / 2
            dist = ((ci[:, None] - cn[None]) ** 2).sum(2)
            track_size = (ib[:, 2] - ib[:, 0]) * (ib[:, 3] - ib[:, 1])
            item_size = (nb[:, 2] - nb[:, 0]) * (nb[:, 3] - nb[:, 1])
            dist = dist + ((dist > track_size[:, None]) | (dist > item_size[None])) * 1e18
            pairs = []
            for i in range(dist.shape[0]):
                j = int(dist[i].argmin())
                if dist[i, j] < 1e16:
                    dist[:, j] = 1e18
                    dist[i, j] = 0.0
                    pairs.append((i, j))
        else:
            from scipy.optimize import linear_sum_assignment
            # embeddings: an inactive track's last decoder output (its row of the query pool), a detection's this frame
            rows_i, rows_n = self._to_device_indices(hs.device, [e[3] for e in inactive], [e[2] for e in new])
            dist = torch.cdist(self.track_hs[rows_i].float(), hs[0, rows_n].float()).cpu().numpy()
            pairs = list(zip(*linear_sum_assignment(dist)))
        taken_i, taken_n = set(), set()
        for i, j in pairs:
            if dist[i, j] <= self.reid_sim_threshold:
                e = new[j]
                live.append([inactive[i][0], 0, e[2], e[2], e[4], e[5]])
                taken_i.add(i); taken_n.add(j)
                self.num_reids += 1
        inactive = [e for i, e in enumerate(inactive) if i not in taken_i]
        new = [e for j, e in enumerate(new) if j not in taken_n]
        return live, inactive, new

    def _update_tracks(self, hs, xyxy, packed, public_dets=None):
        """Host side of a frame: packed = (Lq, 6) numpy [score, label, x1, y1, x2, y2] of all queries."""
        dev = hs.device
        T, I = len(self.track_ids), len(self.inactive_ids)
        sc, lb, bx = packed[:, 0], packed[:, 1].astype(np.int64), packed[:, 2:6].astype(np.float64)

        # tracks (tracker.py:334-420): entry = [id, misses, query index this frame or -1 (stale), old row, box, score]
        live, inactive = [], []                  # (inactive entries carry their count of inactive frames in slot 1)
        for j in range(T):
            if sc[j] > self.trk_thresh and lb[j] == 0:
                live.append([self.track_ids[j], 0, j, j, bx[j], sc[j]])
            elif self.track_misses[j] + 1 < self.steps_termination:      # missed, not yet terminated: keeps its state
                live.append([self.track_ids[j], self.track_misses[j] + 1, -1, j, self._host_boxes[j], self.track_scores[j]])
            elif self.inactive_patience >= 0:                             # terminated: inactive, its last state kept (:88-94)
                inactive.append([self.track_ids[j], 0, -1, j, self._host_boxes[j], self.track_scores[j]])
        # inactive track queries above reid_score_thresh return, behind the live tracks (:367-383, :391-394; the reference
        # does not reset their count of inactive frames on this path)
        stay = []
        for j in range(I):
            r = T + j
            if sc[r] > self.reid_thresh and lb[r] == 0:
                live.append([self.inactive_ids[j], 0, r, r, bx[r], sc[r]])
                self.num_reids += 1
            else:
                stay.append([self.inactive_ids[j], self.inactive_count[j], -1, r, self._host_boxes[r], self.track_scores[r]])
        inactive = stay + inactive               # (move_tracks_to_inactive appends behind the inactive tracks already there)
        T = T + I                                # the object queries follow all track queries
        if self.trk_nms and live:
            k = _nms(np.stack([e[4] for e in live]), np.asarray([e[5] for e in live]), self.trk_nms)
            live = [live[j] for j in k]
        # new detections (tracker.py:422-505)
        obj = slice(T, T + self.num_object_queries)
        new = [[-1, 0, int(i), -1, bx[i], sc[i]]
               for i in T + np.flatnonzero((sc[obj] > self.det_thresh) & (lb[obj] == 0))]
        if self.public_detections and new:                             # :446-448
            ok = public_detections_mask(np.stack([e[4] for e in new]), public_dets, self.public_detections)
            new = [e for e, k in zip(new, ok) if k]
        if inactive and new and (self.reid_greedy or self.reid_sim_threshold > 0):
            live, inactive, new = self._reid(hs, live, inactive, new)
        # NMS in which existing tracks always win over new ones (:485-505)
        if self.det_nms and (live or new):
            boxes_all = np.stack([e[4] for e in live + new])
            scores_all = np.asarray([np.inf] * len(live) + [e[5] for e in new])
            k = set(_nms(boxes_all, scores_all, self.det_nms).tolist())
            new = [e for j, e in enumerate(new) if j + len(live) in k]
            live = [e for j, e in enumerate(live) if j in k]
        for e in new:
            e[0] = self.next_id
            self.next_id += 1
        entries = live + new
        pool = entries + inactive                # next frame's track queries: live tracks, then inactive ones
        # next frame's track state, gathered on the device
        if pool:
            stale = [j for j, e in enumerate(pool) if e[2] < 0]
            q_rows, dst, src_rows = self._to_device_indices(
                dev, [max(e[2], 0) for e in pool], stale, [pool[j][3] for j in stale])
            new_boxes, new_hs = xyxy[0, q_rows], hs[0, q_rows].float()
            if stale:                                               # a missed frame keeps the previous box / embedding
                new_boxes[dst] = self.track_boxes[src_rows]
                new_hs[dst] = self.track_hs[src_rows]
            self.track_boxes, self.track_hs = new_boxes, new_hs
        else:
            self.track_boxes = self.track_hs = None
        self.track_ids = [e[0] for e in entries]
        self.track_misses = [e[1] for e in entries]
        self.inactive_ids = [e[0] for e in inactive]
        self.inactive_count = [e[1] + 1 for e in inactive]           # :549-551
        self._host_boxes = np.asarray([e[4] for e in pool], dtype=np.float64).reshape(-1, 4)
        self.track_scores = np.asarray([e[5] for e in pool], dtype=np.float64)
        frame = {e[0]: {"bbox": np.array(e[4], dtype=np.float64), "score": float(e[5])} for e in entries}
        for t, r in frame.items():
            self.results.setdefault(t, {})[self.frame_index] = r
        self.frame_index += 1
        return frame


def synthetic_sequence(levels, d_model, n_frames, device, seed=0, drift=0.05, dtype=torch.float32):
    """Frames of a synthetic sequence as the transformer sees them: per-level feature maps (1, C, H, W), all-False
    padding masks and positional maps.  Consecutive frames are correlated (a fixed scene plus a slowly varying
    part), so detections persist for a while and then change -- the number of track queries varies from frame to
    frame like it does on video.  Yields (srcs, masks, pos_embeds)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    scene = [torch.randn(1, d_model, h, w, generator=g) for h, w in levels]
    pos = [torch.randn(1, d_model, h, w, generator=g) * 0.1 for h, w in levels]
    masks = [torch.zeros(1, h, w, dtype=torch.bool, device=device) for h, w in levels]
    pos = [p.to(device=device, dtype=dtype) for p in pos]
    cur = [s.clone() for s in scene]
    for _ in range(n_frames):
        cur = [(1 - drift) * c + drift * torch.randn(c.shape, generator=g) + drift * s for c, s in zip(cur, scene)]
        yield [c.to(device=device, dtype=dtype) for c in cur], masks, pos


# =====================================================================================================================
# The literal KineT tracker: kinematic detector over detection tokens
# =====================================================================================================================
def _host_f32(a):
    """numpy float32 view / copy of a host array or tensor."""
    if isinstance(a, torch.Tensor):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


class FrameTracks(Mapping):
    """One frame's tracks as the read-only mapping {track_id: {"bbox": xyxy pixels, "score"}} over three parallel
    arrays (``ids``, ``boxes``, ``scores``); the dicts are made on first access, so a loop that only steps pays nothing."""
    __slots__ = ("ids", "boxes", "scores", "_d")

    def __init__(self, ids, boxes, scores):
        self.ids, self.boxes, self.scores, self._d = ids, boxes, scores, None

    def _dict(self):
        if self._d is None:
            self._d = {int(t): {"bbox": self.boxes[i].copy(), "score": float(self.scores[i])}
                       for i, t in enumerate(self.ids)}
        return self._d

    def __getitem__(self, tid):
        return self._dict()[tid]

    def __iter__(self):
        return iter(self._dict())

    def __len__(self):
        return len(self.ids)

    def __repr__(self):
        return repr(self._dict())


class KineSequenceTracker:
    """Tracks one sequence frame by frame with the `kine` model -- the loop of the reference's ``TrackerKinematic.step``
    (src/trackformer/models/tracker.py:626-880) around ``kinet_b200.modules.kinematic_detector``:

      * :652-671  every live track hands its TRAIL -- its last ``n_frames`` relative boxes and scores (``TrackKinematic.
        get_trail``, :1027-1046; a young track repeats its oldest entry) -- to the detector as a track query
        (``track_query_hs_embeds_det / _meta``; the tracklet encoder is the identity, ``use_encoding_tracklets: False``);
      * :672-690  detector forward, PostProcess (softmax over classes + no-object, best real class; boxes cxcywh -> xyxy
        pixels, clipped unless ``overflow_boxes``);
      * :692-707, :937-955  a track survives while score > ``track_obj_score_thresh`` and its label is class 0 (its trail
        grows by the new box / score), else it counts a miss and ends after ``steps_termination`` misses; NMS among tracks;
      * :750, :895-930  object queries above ``detection_obj_score_thresh`` start tracks whose trail is the detection repeated
        ``n_frames`` times (``generate_pseudo_tracklets``); :820-838 NMS in which existing tracks always win;
      * :860-880  per-frame results.

      * :630-633, :652, :700-713, :922-931  with ``inactive_patience`` >= 0 (cfgs/track_kinet.yaml:45: 5) a terminated track
        stays an INACTIVE track query for that many frames (its trail frozen, the last entry repeated once) and comes back
        under its old id when its query scores above ``reid_score_thresh``; -1 (cfgs/track.yaml) forgets it at once.

    ``public_detections`` ('center_distance' / 'min_iou_0_5', :126-167) gates births like in the reference.  Not covered: the embedding-distance
    ``reid()`` of the deformable Tracker is commented out in TrackerKinematic.step (:756-762).  The reference's
    ``add_tracks`` (:882-893) does not run as shipped (it passes ``confidence=`` to a constructor without that parameter,
    SURVEY.md F7); the evident intent -- a TrackKinematic holding the detection's box, score and pseudo-trail -- is built.

    ``detector``: kinematic_detector.KinematicDetectorTransformer whose backbones accept ``(tokens, pos)`` pairs (e.g. the
    pass-through in tools/kine_bench.py; the reference's Kinet_Backbone + position encoding are the caller's).  A frame is
    ``(src_det, src_meta, pos_det, pos_meta)``, each (1, T, C).  ``use_graphs``: replay kinet_b200.frame_graph.KineFrameGraphs
    (tokens padded to ``token_bucket``, track queries to ``track_bucket``, both masked; one graph per (padded T, padded K)); otherwise the detector runs eagerly through
    its own forward (``targets`` carrying the trails).  One device -> host read per frame; track state lives on the host
    (a few dozen boxes) and goes to the device through pinned staging."""

    def __init__(self, detector, img_size, n_frames=5, detection_obj_score_thresh=0.4, track_obj_score_thresh=0.4,
                 detection_nms_thresh=0.9, track_nms_thresh=0.9, steps_termination=1, n_classes=1, use_graphs=False,
                 token_bucket=32, track_bucket=8, autocast_dtype=None, inactive_patience=-1, reid_score_thresh=0.4,
                 public_detections=False):
        self.detector = detector
        self.public_detections = public_detections       # False | 'center_distance' | 'min_iou_0_5' (:126-167, :905-910)
        self.inactive_patience, self.reid_thresh = inactive_patience, reid_score_thresh
        self.img_h, self.img_w = float(img_size[0]), float(img_size[1])
        self._wh = np.array([self.img_w, self.img_h, self.img_w, self.img_h])
        self.n_frames = int(n_frames)
        self.det_thresh, self.trk_thresh = detection_obj_score_thresh, track_obj_score_thresh
        self.det_nms, self.trk_nms = detection_nms_thresh, track_nms_thresh
        self.steps_termination, self.n_classes = steps_termination, n_classes
        self.use_graphs, self.token_bucket, self.autocast_dtype = use_graphs, token_bucket, autocast_dtype
        self.track_bucket = track_bucket
        self._graphs = None
        self.reset()

    def reset(self):
        n = self.n_frames
        # host-side track state, one row per live track (a few dozen): struct of arrays, so a frame costs O(1) numpy calls
        self.ids = np.empty(0, np.int64)
        self.boxes = np.empty((0, 4), np.float64)            # xyxy pixels
        self.scores = np.empty(0, np.float64)
        self.misses = np.empty(0, np.int64)
        self.trail = np.empty((0, n, 4), np.float32)         # last n relative boxes (cx, cy, w, h), oldest first
        self.trail_sc = np.empty((0, n), np.float32)         # ... and scores
        # terminated tracks kept as track queries for ``inactive_patience`` more frames (re-identification, :630-633):
        # ids, boxes, scores, trail, trail scores, frames spent inactive
        self.inactive = (np.empty(0, np.int64), np.empty((0, 4), np.float64), np.empty(0, np.float64),
                         np.empty((0, n, 4), np.float32), np.empty((0, n), np.float32), np.empty(0, np.int64))
        self.num_reids = 0
        self.next_id = 0
        self.frame_index = 0
        self._frames = []

    @property
    def tracks(self):
        return self.ids.tolist()

    @property
    def inactive_tracks(self):
        return self.inactive[0].tolist()

    def _prune_inactive(self):
        """Start of a frame (:630-633): inactive tracks beyond their patience, or without area, are forgotten."""
        iid, ibox, isc, itr, itsc, icnt = self.inactive
        if len(iid):
            ok = (icnt <= self.inactive_patience) & (ibox[:, 2] > ibox[:, 0]) & (ibox[:, 3] > ibox[:, 1])
            if not ok.all():
                self.inactive = tuple(a[ok] for a in self.inactive)

    @property
    def results(self):
        """{track_id: {frame_index: {"bbox", "score"}}} (the reference's ``get_results``, tracker.py:568-573), built
        from the per-frame arrays when asked for: the loop itself stores three array references per frame."""
        res = {}
        for f, frame in enumerate(self._frames):
            for tid, r in frame.items():
                res.setdefault(tid, {})[f] = r
        return res

    # ---- trails -> track-query inputs (tracker.py:652-671, :1027-1046) ----
    def _trails(self):
        """Track queries of the frame: the live tracks, then the inactive ones (``self.tracks + self.inactive_tracks``, :652)."""
        trail, trail_sc = self.trail, self.trail_sc
        if len(self.inactive[0]):
            trail, trail_sc = np.concatenate([trail, self.inactive[3]]), np.concatenate([trail_sc, self.inactive[4]])
        return trail.reshape(len(trail), 4 * self.n_frames), trail_sc

    def _to_device(self, dev, det, meta):
        """(K, n*4) and (K, n) host arrays -> device tensors through pinned staging (see SequenceTracker._to_device_indices)."""
        n = det.size + meta.size
        st = getattr(self, "_stage", None)
        if st is None or st[0].shape[0] < n:
            cap = max(4096, 2 * n)
            st = self._stage = [torch.empty(cap).pin_memory(), torch.empty(cap).pin_memory(), 0]
        st[2] ^= 1
        host = st[st[2]]
        flat = host.numpy()
        flat[:det.size] = det.ravel()
        flat[det.size:n] = meta.ravel()
        d = host[:n].to(dev, non_blocking=True)
        return d[:det.size].view(det.shape), d[det.size:].view(meta.shape)

    # ---- one detector call -> packed (Lq, 6): score, label, box cx cy w h (relative) ----
    def _pack(self, logits, boxes):
        prob = logits.float().softmax(-1)[..., :-1]                     # PostProcess: no-object is the last class
        scores, labels = prob.max(-1)
        return torch.cat([scores[0, :, None], labels[0, :, None].to(scores.dtype), boxes[0].float()], -1)

    def _detect_eager(self, frame, det_t, meta_t):

        class _S:
            pass
        s = _S()
        if len(frame) == 2:                       # raw detections (T, Xb) / metadata (T, Xc): through the model's backbones
            from .modules.kinematic_backbone import NestedTensor
            dev = self.query_device()
            d, m = (torch.from_numpy(_host_f32(a)).to(dev)[None] for a in frame)
            mask = torch.zeros(1, d.shape[1], dtype=torch.bool, device=dev)
            s.detections, s.metadata = NestedTensor(d, mask), NestedTensor(m, mask)
        else:
            s.detections, s.metadata = (frame[0], frame[2]), (frame[1], frame[3])
        targets = None
        if det_t is not None:
            targets = [{"track_query_hs_embeds_det": det_t, "track_query_hs_embeds_meta": meta_t}]
        with torch.autocast("cuda", dtype=self.autocast_dtype, enabled=self.autocast_dtype is not None):
            out = self.detector(s, targets)[0]
        return self._pack(out["pred_logits"], out["pred_boxes"])

    def _detect_graph(self, frame, trail_det, trail_meta):
        """trail_det (K, n*4) / trail_meta (K, n) HOST arrays: KineFrameGraphs.call_with_trails moves them (and the
        frame's counts) in one pinned block; the trail projections are part of the recorded graph."""
        from .frame_graph import KineFrameGraphs
        from .functions.layer_ops import linear
        det = self.detector
        if self._graphs is None:
            nq = det.num_queries

            def heads(transformer, bufs, Kp):
                pad = lambda e: torch.nn.functional.pad(e.weight[:, None], (0, 0, 0, 0, Kp, 0))
                if "det_raw" in bufs:              # a frame from its raw detections: the backbones are part of the graph
                    from .modules.kinematic_backbone import NestedTensor
                    fd, pd = det.backbone_det(NestedTensor(bufs["det_raw"], bufs["mask"]))
                    fm, pm = det.backbone_metadata(NestedTensor(bufs["meta_raw"], bufs["mask"]))
                    src_det, src_meta, pos_det, pos_meta = fd[-1].tensors, fm[-1].tensors, pd[0], pm[0]
                else:
                    src_det, src_meta, pos_det, pos_meta = bufs["src_det"], bufs["src_meta"], bufs["pos_det"], bufs["pos_meta"]
                td = tm = qpad = None
                if Kp:
                    td = det.input_proj_tracklets_det(bufs["trail_det"])[:, None].float()
                    tm = det.input_proj_tracklets_metadata(bufs["trail_meta"])[:, None].float()
                    td = torch.nn.functional.pad(td, (0, 0, 0, 0, 0, nq))
                    tm = torch.nn.functional.pad(tm, (0, 0, 0, 0, 0, nq))
                    qpad = torch.nn.functional.pad(bufs["kpad"], (0, nq))                 # object queries are never padding
                hs_det, hs_meta, _, _ = transformer.forward_masked_queries(
                    src_det, src_meta, bufs["mask"], pad(det.query_embed_det), pad(det.query_embed_metadata),
                    td, tm, pos_det, pos_meta, qpad)
                return self._pack(linear(hs_meta[-1], det.class_embed), det.bbox_embed(hs_det[-1]).sigmoid())
            self._graphs = KineFrameGraphs(det.transformer, heads, token_bucket=self.token_bucket,
                                           track_bucket=self.track_bucket, autocast_dtype=self.autocast_dtype)
        if len(frame) == 2:
            dev = self.query_device()
            packed, pad = self._graphs.call_with_detections(dev, _host_f32(frame[0]), _host_f32(frame[1]), trail_det, trail_meta)
        else:
            packed, pad = self._graphs.call_with_trails(frame[0], frame[1], frame[2], frame[3], trail_det, trail_meta)
        return packed[pad:]                       # the padding track-query rows sit at the front

    def query_device(self):
        return self.detector.query_embed_det.weight.device

    @torch.no_grad()
    def step(self, frame, public_dets=None):
        """frame = (src_det, src_meta, pos_det, pos_meta), each (1, T, C) on the device -- or, with a detector built on
        modules.kinematic_backbone, the frame's RAW detections (boxes (T, 4), metadata (T, 1 or 2)) as host arrays
        -> {track_id: {"bbox", "score"}} (a read-only mapping built on first access; ``.ids`` / ``.boxes`` / ``.scores``
        are the arrays behind it)."""
        dev = self.query_device()
        self._prune_inactive()
        K = len(self.ids) + len(self.inactive[0])
        if self.use_graphs:
            packed = self._detect_graph(frame, *self._trails())
        else:
            det_t = meta_t = None
            if K:
                det_t, meta_t = self._to_device(dev, *self._trails())
            packed = self._detect_eager(frame, det_t, meta_t)
        return self._update(packed.cpu().numpy(), K, public_dets)         # ONE device -> host read per frame

    def _update(self, packed, K, public_dets=None):
        """packed (K + num_queries, 6) float32 from the detector (score, label, relative cx cy w h), the K track queries
        (live tracks, then inactive ones) first.  Every state array is REPLACED (FrameTracks of earlier frames keep theirs);
        selects are whole-array ``np.where`` / boolean compress, and both suppression passes read one IoU matrix."""
        n = self.n_frames
        nA = len(self.ids)                                                 # live tracks; rows nA .. K-1 are inactive ones
        p = packed.astype(np.float64)
        sc, lb = p[:, 0], p[:, 1]
        rel = np.clip(p[:, 2:6], 0.0, 1.0)                                 # relative boxes are clamped (:688-690)
        half = rel[:, 2:] * 0.5
        px = np.concatenate([rel[:, :2] - half, rel[:, :2] + half], 1)
        px *= self._wh
        np.clip(px, 0.0, self._wh, out=px)                                 # clip_boxes_to_image
        rel32 = rel.astype(np.float32)
        ids, boxes, scores, misses, trail, trail_sc = self.ids, self.boxes, self.scores, self.misses, self.trail, self.trail_sc
        gone = None                                                        # live tracks terminated in this frame
        if nA:                                                             # manage_active_tracks (:937-955)
            keep = (sc[:nA] > self.trk_thresh) & (lb[:nA] == 0)
            k1, k2 = keep[:, None], keep[:, None, None]
            boxes = np.where(k1, px[:nA], boxes)
            scores = np.where(keep, sc[:nA], scores)
            misses = np.where(keep, 0, misses + 1)
            # a surviving track's trail grows by the new box / score (young tracks repeat their oldest entry: a trail that
            # starts as n copies of the first detection, shifted left once per frame, is exactly TrackKinematic.get_trail)
            trail = np.where(k2, np.concatenate([trail[:, 1:], rel32[:nA, None]], 1), trail)
            trail_sc = np.where(k1, np.concatenate([trail_sc[:, 1:], sc[:nA, None].astype(np.float32)], 1), trail_sc)
            alive = keep | (misses < self.steps_termination)
            if not alive.all():
                if self.inactive_patience >= 0:
                    gone = tuple(a[~alive] for a in (ids, boxes, scores, trail, trail_sc))
                ids, boxes, scores, misses, trail, trail_sc = (a[alive] for a in (ids, boxes, scores, misses, trail, trail_sc))
        if K > nA or gone is not None:
            # inactive track queries that score above reid_score_thresh come back (manage_inactive_tracks, :922-931: behind
            # the live tracks; the reference does not reset their count of inactive frames, :241 is the other re-id path);
            # the tracks terminated above go inactive with their last state repeated once (move_tracks_to_inactive, :88-94,
            # whose repeat_last_pos() is TrackKinematic.repeat_last_state -- the name it calls does not exist)
            iid, ibox, isc, itr, itsc, icnt = self.inactive
            if K > nA:
                back = (sc[nA:K] > self.reid_thresh) & (lb[nA:K] == 0)
                if back.any():
                    self.num_reids += int(back.sum())
                    ids = np.concatenate([ids, iid[back]])
                    boxes = np.concatenate([boxes, px[nA:K][back]])
                    scores = np.concatenate([scores, sc[nA:K][back]])
                    misses = np.concatenate([misses, np.zeros(int(back.sum()), np.int64)])
                    trail = np.concatenate([trail, np.concatenate([itr[back][:, 1:], rel32[nA:K][back][:, None]], 1)])
                    trail_sc = np.concatenate([trail_sc, np.concatenate([itsc[back][:, 1:], sc[nA:K][back][:, None].astype(np.float32)], 1)])
                    iid, ibox, isc, itr, itsc, icnt = (a[~back] for a in (iid, ibox, isc, itr, itsc, icnt))
            if gone is not None:
                gid, gbox, gsc, gtr, gtsc = gone
                iid, ibox, isc = np.concatenate([iid, gid]), np.concatenate([ibox, gbox]), np.concatenate([isc, gsc])
                itr = np.concatenate([itr, np.concatenate([gtr[:, 1:], gtr[:, -1:]], 1)])
                itsc = np.concatenate([itsc, np.concatenate([gtsc[:, 1:], gtsc[:, -1:]], 1)])
                icnt = np.concatenate([icnt, np.zeros(len(gid), np.int64)])
            self.inactive = (iid, ibox, isc, itr, itsc, icnt)
        # generate_new_tracks (:895-930): object queries above the detection threshold
        idx = K + np.flatnonzero((sc[K:] > self.det_thresh) & (lb[K:] < self.n_classes))
        if self.public_detections and len(idx):
            idx = idx[public_detections_mask(px[idx], public_dets, self.public_detections)]
        nl, m = len(ids), len(idx)
        sel_t, sel_n = None, None                                          # None = all kept
        need_t, need_d = bool(self.trk_nms) and nl > 1, bool(self.det_nms) and nl + m > 1
        if need_t or need_d:
            allb = np.concatenate([boxes, px[idx]]) if m else boxes
            iou = _iou_matrix(allb)
            iou.flat[::nl + m + 1] = 0.0
            lo = min(t for t, need in ((self.trk_nms, need_t), (self.det_nms, need_d)) if need)
            if (iou > lo).any():                                           # rare: something overlaps, run the greedy passes
                kt, kn = np.arange(nl), np.arange(m)
                if need_t:                                                 # NMS among the tracks (:733)
                    kt = _greedy_keep(iou[:nl, :nl] > self.trk_nms, np.argsort(-scores, kind="stable"))
                if need_d:                                                 # :786-800: tracks, old ones at score inf, and births
                    sel = np.concatenate([kt, nl + kn])
                    order = np.concatenate([np.arange(len(kt)), len(kt) + np.argsort(-sc[idx], kind="stable")])
                    k = _greedy_keep(iou[np.ix_(sel, sel)] > self.det_nms, order)
                    kt, kn = kt[k[k < len(kt)]], k[k >= len(kt)] - len(kt)
                sel_t, sel_n = kt, kn
        if sel_t is not None and len(sel_t) < nl:
            ids, boxes, scores, misses, trail, trail_sc = (a[sel_t] for a in (ids, boxes, scores, misses, trail, trail_sc))
        if sel_n is not None and len(sel_n) < m:
            idx = idx[sel_n]
            m = len(idx)
        if m:
            ns = sc[idx]
            self.ids = np.concatenate([ids, np.arange(self.next_id, self.next_id + m, dtype=np.int64)])
            self.next_id += m
            self.boxes = np.concatenate([boxes, px[idx]])
            self.scores = np.concatenate([scores, ns])
            self.misses = np.concatenate([misses, np.zeros(m, np.int64)])
            # the detection repeated n times is the new track's trail (generate_pseudo_tracklets)
            self.trail = np.concatenate([trail, np.repeat(rel32[idx][:, None], n, 1)])
            self.trail_sc = np.concatenate([trail_sc, np.repeat(ns[:, None].astype(np.float32), n, 1)])
        else:
            self.ids, self.boxes, self.scores, self.misses, self.trail, self.trail_sc = ids, boxes, scores, misses, trail, trail_sc
        if len(self.inactive[0]):                                          # :849-850
            self.inactive = self.inactive[:5] + (self.inactive[5] + 1,)
        frame = FrameTracks(self.ids, self.boxes, self.scores)             # the arrays are replaced, never edited in place
        self._frames.append(frame)
        self.frame_index += 1
        return frame
