#!/usr/bin/env python
"""Frames/s of the literal KineT (`kine`) transformer path at the reference's configuration (cfgs/train_kinet.yaml:
hidden_dim 288, 8 heads, 1 encoder + 1 decoder layer, 150 object queries; dim_feedforward 2048, cfgs/train.yaml:51):

  fused    kinet_b200.modules.kinematic_transformer.DualKinematicTransformer + the detector heads, recorded into
           per-(tokens, track queries) CUDA graphs (kinet_b200.frame_graph.KineFrameGraphs)
  plain    the same parameters through nn.MultiheadAttention / nn.LayerNorm / nn.Linear written as the reference's
           forward_post does (src/trackformer/models/transformer.py:323-339, :385-407, :483-487), eager and as one
           CUDA graph

Synthetic detection tokens (T ~ 100-150 per frame, changing every frame) and K ~ 20-60 track queries; bf16 autocast
(--fp32 for fp32).  Prints one JSON line.
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(dev, d=288, nq=150):
    from kinet_b200.modules.kinematic_detector import MLP
    from kinet_b200.modules.kinematic_transformer import DualKinematicTransformer
    torch.manual_seed(0)
    tr = DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1, dim_feedforward=2048,
                                  dropout=0.1, return_intermediate_dec=True).to(dev).eval()
    mods = torch.nn.ModuleDict(dict(class_embed=torch.nn.Linear(d, 2), bbox_embed=MLP(d, d, 4, 3),
                                    query_embed_det=torch.nn.Embedding(nq, d),
                                    query_embed_metadata=torch.nn.Embedding(nq, d))).to(dev).eval()
    return tr, mods


def make_heads(mods):
    from kinet_b200.functions.layer_ops import linear

    def heads(transformer, bufs, Kp):
        qd = F.pad(mods["query_embed_det"].weight[:, None], (0, 0, 0, 0, Kp, 0))
        qm = F.pad(mods["query_embed_metadata"].weight[:, None], (0, 0, 0, 0, Kp, 0))
        nq = mods["query_embed_det"].weight.shape[0]
        td = F.pad(bufs["tgt_det"], (0, 0, 0, 0, 0, nq)) if Kp else None
        tm = F.pad(bufs["tgt_meta"], (0, 0, 0, 0, 0, nq)) if Kp else None
        qpad = F.pad(bufs["kpad"], (0, nq)) if Kp else None
        hs_det, hs_meta, _, _ = transformer.forward_masked_queries(bufs["src_det"], bufs["src_meta"], bufs["mask"], qd, qm,
                                                                   td, tm, bufs["pos_det"], bufs["pos_meta"], qpad)
        return linear(hs_meta[-1], mods["class_embed"]).float(), mods["bbox_embed"](hs_det[-1]).float().sigmoid(), hs_det[-1]
    return heads


def plain_forward(tr, mods, src_b, src_m, mask, qd, qm, td, tm, pos_b, pos_m, p_drop=0.0):
    """The reference's arithmetic on the same parameters with stock PyTorch modules (sequence-first inside).
    p_drop > 0: the residual-branch and FFN dropouts of forward_post (training)."""
    dp = (lambda x: F.dropout(x, p_drop, True)) if p_drop > 0 else (lambda x: x)

    def enc_layer(l, src, pos):
        q = src + pos
        src = l.norm1(src + dp(l.self_attn(q, q, value=src, key_padding_mask=mask, need_weights=False)[0]))
        return l.norm2(src + dp(l.linear2(dp(F.relu(l.linear1(src))))))

    def dec_layer(l, tgt, mem, pos, qpos):
        q = tgt + qpos
        tgt = l.norm1(tgt + dp(l.self_attn(q, q, value=tgt, need_weights=False)[0]))
        tgt = l.norm2(tgt + dp(l.multihead_attn(tgt + qpos, mem + pos, value=mem, key_padding_mask=mask, need_weights=False)[0]))
        return l.norm3(tgt + dp(l.linear2(dp(F.relu(l.linear1(tgt))))))

    def one(t, src, pos, qe, tgt):
        src, pos = src.permute(1, 0, 2), pos.permute(1, 0, 2)
        tgt = torch.zeros_like(qe) if tgt is None else tgt
        for l in t.encoder.layers:
            src = enc_layer(l, src, pos)
        inter = []
        for l in t.decoder.layers:
            tgt = dec_layer(l, tgt, src, pos, qe)
            inter.append(tgt)
        return t.decoder.norm(torch.stack(inter))

    def branch(b, s1, s2):
        return b.norm(F.relu(torch.cat([b.linear_input1(s1), b.linear_input2(s2)], -1)) + s1)
    hd = one(tr.transformer_det, src_b, pos_b, qd, td)
    hm = one(tr.transformer_metadata, src_m, pos_m, qm, tm)
    hd = branch(tr.detection_branch, hd, hm)
    hm = branch(tr.metadata_branch, hm, hd)
    hd, hm = hd.transpose(1, 2), hm.transpose(1, 2)
    x = hd[-1]
    for i, layer in enumerate(mods["bbox_embed"].layers):
        x = F.relu(layer(x)) if i < 2 else layer(x)
    return mods["class_embed"](hm[-1]).float(), x.float().sigmoid(), hd[-1]


def measure(dev, frames=200, fp32=False, nq=150, d=288, seed=0):
    from kinet_b200.frame_graph import KineFrameGraphs
    tr, mods = build(dev, d, nq)
    amp = None if fp32 else torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(seed)
    sizes = [(100 + (37 * i) % 51, 20 + (13 * i) % 41) for i in range(16)]        # (tokens, track queries) per frame
    feeds = []
    for T, K in sizes:
        feeds.append(tuple(torch.randn(1, T, d, device=dev, generator=g) for _ in range(4)) +
                     tuple(torch.randn(K, 1, d, device=dev, generator=g) for _ in range(2)))
    fg = KineFrameGraphs(tr, make_heads(mods), token_bucket=32, track_bucket=16, autocast_dtype=amp)
    for f in feeds:
        fg(*f)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(frames):
        out, pad = fg(*feeds[i & 15])
    b.record()
    torch.cuda.synchronize()
    fused_ms = a.elapsed_time(b) / frames

    # plain PyTorch modules on the same parameters: eager (what the reference runs) and one graph at a fixed size
    T, K = 150, 48
    src_b, src_m, pos_b, pos_m = (torch.randn(1, T, d, device=dev, generator=g) for _ in range(4))
    mask = torch.zeros(1, T, dtype=torch.bool, device=dev)
    qd = F.pad(mods["query_embed_det"].weight[:, None], (0, 0, 0, 0, K, 0)).detach()
    qm = F.pad(mods["query_embed_metadata"].weight[:, None], (0, 0, 0, 0, K, 0)).detach()
    td = F.pad(torch.randn(K, 1, d, device=dev, generator=g), (0, 0, 0, 0, 0, nq))
    tm = F.pad(torch.randn(K, 1, d, device=dev, generator=g), (0, 0, 0, 0, 0, nq))

    def plain():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=not fp32):
            return plain_forward(tr, mods, src_b, src_m, mask, qd, qm, td, tm, pos_b, pos_m)

    def fused_fixed():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=not fp32):
            hd, hm, _, _ = tr(src_b, src_m, mask, qd, qm, td, tm, pos_b, pos_m)
            return hd

    def timed(fn, n):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n

    def graphed(fn):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn(); fn()
        torch.cuda.current_stream().wait_stream(s)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        return gr.replay
    with torch.no_grad():
        want = plain()[2]
        got = fused_fixed()[-1]
    err = float((got.float() - want.float()).abs().max() / want.float().abs().max())
    res = dict(fused_graphs_ms_per_frame=fused_ms, fused_frames_per_s=1e3 / fused_ms,
               graphs_recorded=len(fg.recorded_buckets),
               plain_eager_ms=timed(plain, 50), plain_graph_ms=timed(graphed(plain), 200),
               fused_eager_ms=timed(fused_fixed, 50), fused_graph_fixed_ms=timed(graphed(fused_fixed), 200),
               fused_vs_plain_max_rel=err, dtype="fp32" if fp32 else "bf16 autocast",
               config="DualKinematicTransformer d=288, 8 heads (D=36), 1+1 layers, ffn 2048, %d object queries + K "
                      "track queries (20-60), T detection tokens (100-150), batch 1" % nq)
    return res


def measure_tracker(dev, frames=200, fp32=False, nq=150, d=288, n_frames=5, seed=0, raw=False):
    """Frames/s of kinet_b200.tracking.KineSequenceTracker (the loop of TrackerKinematic.step) over the kine detector at the
    kinet configuration: per frame T detection tokens (100-150), the live tracks' trails as track queries, one detector
    replay from KineFrameGraphs, ONE read-back, host thresholds / NMS / births / terminations.
    raw: the frames are RAW detections (boxes + confidences, host arrays) and the detector carries the model's own
    backbones (modules.kinematic_backbone): the frame is one pinned block + one replay."""
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer, TokenFeatures
    from kinet_b200.modules.kinematic_transformer import DualKinematicTransformer
    from kinet_b200.tracking import KineSequenceTracker
    torch.manual_seed(seed)
    back = [lambda x: ([TokenFeatures(x[0], torch.zeros(x[0].shape[:2], dtype=torch.bool, device=x[0].device))], [x[1]])] * 2
    if raw:
        from kinet_b200.modules.kinematic_backbone import build_kine_backbone
        back = build_kine_backbone(d, max_number_detection=200)
    tr = DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1, dim_feedforward=2048,
                                  dropout=0.1, return_intermediate_dec=True)
    det = KinematicDetectorTransformer(back, tr, num_classes=1, num_queries=nq, dim_tracklets_det=4 * n_frames,
                                       dim_tracklets_metadata=n_frames).to(dev).eval()
    with torch.no_grad():                       # random weights: spread scores and boxes so that tracks come and go
        det.class_embed.weight.mul_(4.0)
        det.bbox_embed.layers[-1].weight.mul_(8.0)
    g = torch.Generator(device=dev).manual_seed(seed + 1)
    scene = [torch.randn(1, 160, d, device=dev, generator=g) for _ in range(4)]
    feeds = []
    import numpy as np
    rng = np.random.default_rng(seed + 2)
    boxes = rng.random((160, 5)).astype(np.float32)
    for i in range(16):
        T = 100 + (37 * i) % 51
        if raw:
            f = np.clip(boxes[:T] + 0.01 * rng.standard_normal((T, 5)).astype(np.float32), 0, 1)
            feeds.append((np.ascontiguousarray(f[:, :4]), np.ascontiguousarray(f[:, 4:])))
        else:
            feeds.append(tuple(s_[:, :T] + 0.3 * torch.randn(1, T, d, device=dev, generator=g) for s_ in scene))
    trk = KineSequenceTracker(det, img_size=(1080, 1920), n_frames=n_frames, use_graphs=True, token_bucket=32,
                              autocast_dtype=None if fp32 else torch.bfloat16)
    with torch.no_grad():
        sc = trk._detect_eager(feeds[0], None, None)[:, 0].float()
    trk.det_thresh = float(torch.quantile(sc, 0.85))            # ~ 20 of the 150 object queries fire on the first frame
    trk.trk_thresh = trk.det_thresh - 0.05 * float(sc.max() - sc.min())
    if raw:
        # random weights: what a track query scores has little to do with what started it.  Pilot frames in which no track
        # ends give the distribution of track-query scores; the track threshold is put at its 30 % quantile and the
        # detection threshold where ~12 object queries fire, i.e. ~40 live tracks in the steady state.
        import numpy as np
        seen, orig = [], trk._update

        def spy(packed, K, *rest):
            if K:
                seen.append(packed[:K, 0].copy())
            return orig(packed, K, *rest)
        trk._update, trk.trk_thresh, trk.det_thresh = spy, float("-inf"), float(torch.quantile(sc, 0.92))
        for i in range(6):
            trk.step(feeds[i & 15])
        trk._update = orig
        trk.trk_thresh = float(np.quantile(np.concatenate(seen), 0.3)) if seen else trk.det_thresh
        trk.reset()
    for i in range(48):                                         # warm-up: records the graphs of the sizes that occur
        trk.step(feeds[i & 15])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    counts = []
    a.record()
    for i in range(frames):
        counts.append(len(trk.step(feeds[i & 15])))
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / frames
    # where a frame goes (host clock, a second pass): enqueue (trails -> pinned block, copies, replay) / wait for the
    # read-back / host update
    import time
    ph = [0.0, 0.0, 0.0]
    for i in range(frames):
        t0 = time.perf_counter()
        K = len(trk.ids)
        packed = trk._detect_graph(feeds[i & 15], *trk._trails())
        t1 = time.perf_counter()
        host = packed.cpu().numpy()
        t2 = time.perf_counter()
        trk._update(host, K)
        t3 = time.perf_counter()
        ph[0] += t1 - t0; ph[1] += t2 - t1; ph[2] += t3 - t2
    return dict(tracker_ms_per_frame=ms, tracker_frames_per_s=1e3 / ms, tracks_min_max=[min(counts), max(counts)],
                tracks_started=trk.next_id, graphs_recorded=len(trk._graphs.recorded_buckets),
                phases_ms=dict(enqueue=1e3 * ph[0] / frames, wait_and_read_back=1e3 * ph[1] / frames,
                               host_update=1e3 * ph[2] / frames))


def measure_training(dev, steps=30, fp32=False, batch=8, T=150, nq=150, d=288):
    """One training step of the `kine` transformer + heads at the reference's configuration (cfgs/train_kinet.yaml:
    batch_size 8, hidden_dim 288, 1 + 1 layers, 150 queries, dropout 0.1; AdamW lr 1e-4, clip_max_norm 0.1) on synthetic
    token embeddings and a synthetic loss (cross-entropy on the class logits + L1 on the boxes; matcher and SetCriterion
    are out of scope): kinet_b200's modules in kinet_b200.training.GraphedTrainStep (forward + backward + clip + AdamW as
    two CUDA graphs) and eagerly, against the same parameters through stock PyTorch modules (eager, fused AdamW)."""
    from kinet_b200.functions.layer_ops import linear
    from kinet_b200.training import GraphedTrainStep
    amp = dict(device_type="cuda", dtype=torch.bfloat16, enabled=not fp32)
    g = torch.Generator(device=dev).manual_seed(3)
    src_b, src_m, pos_b, pos_m = (torch.randn(batch, T, d, device=dev, generator=g) for _ in range(4))
    mask = torch.zeros(batch, T, dtype=torch.bool, device=dev)
    for i in range(batch):
        mask[i, T - 5 * i:] = True                              # samples with 150, 145, ... detections
    tgt_cls = torch.randint(0, 2, (batch * nq,), device=dev, generator=g)
    tgt_box = torch.rand(batch, nq, 4, device=dev, generator=g)

    def loss_of(logits, boxes):
        return F.cross_entropy(logits.float().flatten(0, 1), tgt_cls) + 5.0 * F.l1_loss(boxes.float(), tgt_box)

    def make(train_mode=True):
        tr, mods = build(dev, d, nq)
        net = torch.nn.ModuleList([tr, mods])
        net.train(train_mode)
        return tr, mods, net

    def queries(mods):
        return (mods["query_embed_det"].weight[:, None].expand(-1, batch, -1),
                mods["query_embed_metadata"].weight[:, None].expand(-1, batch, -1))

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n

    # ---- kinet_b200 modules, two CUDA graphs per step ----
    tr, mods, net = make()

    def fused_loss(src_b, src_m, pos_b, pos_m):
        with torch.autocast(**amp):
            qd, qm = queries(mods)
            hs_det, hs_meta, _, _ = tr(src_b, src_m, mask, qd, qm, None, None, pos_b, pos_m)
            return loss_of(linear(hs_meta[-1], mods["class_embed"]), mods["bbox_embed"](hs_det[-1]).sigmoid())
    step = GraphedTrainStep(net, fused_loss, lambda p: torch.optim.AdamW(p, lr=1e-4, weight_decay=1e-4, fused=True, capturable=True),
                            (src_b, src_m, pos_b, pos_m), clip_max_norm=0.1,
                            lp_dtype=torch.float32 if fp32 else torch.bfloat16)
    graph_ms = timed(lambda: step(src_b, src_m, pos_b, pos_m), steps)
    loss_graph = float(step.loss)

    # ---- kinet_b200 modules, eager ----
    tr, mods, net = make()
    opt = torch.optim.AdamW(net.parameters(), lr=1e-4, weight_decay=1e-4, fused=True)

    def fused_eager():
        opt.zero_grad(set_to_none=True)
        fused_loss_eager = None
        with torch.autocast(**amp):
            qd, qm = queries(mods)
            hs_det, hs_meta, _, _ = tr(src_b, src_m, mask, qd, qm, None, None, pos_b, pos_m)
            fused_loss_eager = loss_of(linear(hs_meta[-1], mods["class_embed"]), mods["bbox_embed"](hs_det[-1]).sigmoid())
        fused_loss_eager.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 0.1)
        opt.step()
    eager_ms = timed(fused_eager, steps)
    if os.environ.get("KINE_TRAIN_PROFILE"):                   # kernel breakdown of one fused training step (stderr)
        from torch.profiler import ProfilerActivity, profile as prof
        with prof(activities=[ProfilerActivity.CUDA]) as pr:
            fused_eager()
            torch.cuda.synchronize()
        print(pr.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=80), file=sys.stderr)

    # ---- the same parameters through stock PyTorch modules (dropout of the residual branches as the reference applies it) ----
    tr2, mods2, net2 = make()
    opt2 = torch.optim.AdamW(net2.parameters(), lr=1e-4, weight_decay=1e-4, fused=True)

    def plain_eager():
        opt2.zero_grad(set_to_none=True)
        with torch.autocast(**amp):
            qd, qm = queries(mods2)
            logits, boxes, _ = plain_forward(tr2, mods2, src_b, src_m, mask, qd, qm, None, None, pos_b, pos_m, p_drop=0.1)
            loss = loss_of(logits, boxes)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net2.parameters(), 0.1)
        opt2.step()
    plain_ms = timed(plain_eager, steps)

    # ---- ... and recorded as ONE CUDA graph per step (PyTorch's whole-network capture recipe, capturable AdamW) ----
    tr3, mods3, net3 = make()
    opt3 = torch.optim.AdamW(net3.parameters(), lr=1e-4, weight_decay=1e-4, fused=True, capturable=True)

    def plain_step():
        with torch.autocast(**amp):
            qd, qm = queries(mods3)
            logits, boxes, _ = plain_forward(tr3, mods3, src_b, src_m, mask, qd, qm, None, None, pos_b, pos_m, p_drop=0.1)
            loss = loss_of(logits, boxes)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net3.parameters(), 0.1)
        opt3.step()
    plain_graph_ms = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                opt3.zero_grad(set_to_none=True)
                plain_step()
        torch.cuda.current_stream().wait_stream(side)
        gr = torch.cuda.CUDAGraph()
        opt3.zero_grad(set_to_none=True)
        with torch.cuda.graph(gr):
            plain_step()
        plain_graph_ms = timed(gr.replay, steps)
    except Exception as e:                                    # noqa: BLE001  (reported, not fatal: it is the comparison arm)
        plain_graph_ms = "capture failed: %s" % str(e)[:120]
    return dict(train_graph_ms_per_step=graph_ms, train_samples_per_s=batch * 1e3 / graph_ms, train_fused_eager_ms=eager_ms,
                train_plain_eager_ms=plain_ms, train_plain_graph_ms=plain_graph_ms, loss_after=loss_graph, batch=batch, tokens=T,
                dtype="fp32" if fp32 else "bf16 autocast")


def profile(dev, fp32):
    """Kernel breakdown (torch profiler, eager) of one fused and one plain frame at T = 150, K = 48."""
    from torch.profiler import ProfilerActivity, profile as prof
    d, nq, T, K = 288, 150, 150, 48
    tr, mods = build(dev, d, nq)
    g = torch.Generator(device=dev).manual_seed(0)
    src_b, src_m, pos_b, pos_m = (torch.randn(1, T, d, device=dev, generator=g) for _ in range(4))
    mask = torch.zeros(1, T, dtype=torch.bool, device=dev)
    qd = F.pad(mods["query_embed_det"].weight[:, None], (0, 0, 0, 0, K, 0)).detach()
    td = F.pad(torch.randn(K, 1, d, device=dev, generator=g), (0, 0, 0, 0, 0, nq))
    for name, fn in (("fused", lambda: tr(src_b, src_m, mask, qd, qd, td, td, pos_b, pos_m)),
                     ("plain", lambda: plain_forward(tr, mods, src_b, src_m, mask, qd, qd, td, td, pos_b, pos_m))):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=not fp32):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            with prof(activities=[ProfilerActivity.CUDA]) as pr:
                for _ in range(10):
                    fn()
                torch.cuda.synchronize()
        print("==", name, "fp32" if fp32 else "bf16", "(10 frames)")
        print(pr.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=70))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=200)
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--profile", action="store_true")
    a = ap.parse_args()
    if a.profile:
        profile(torch.device("cuda:0"), a.fp32)
        sys.exit(0)
    res = measure(torch.device("cuda:0"), a.frames, a.fp32)
    res["tracker"] = measure_tracker(torch.device("cuda:0"), a.frames, a.fp32)
    res["tracker_from_raw_detections"] = measure_tracker(torch.device("cuda:0"), a.frames, a.fp32, raw=True)
    res["training"] = measure_training(torch.device("cuda:0"), fp32=a.fp32)
    print(json.dumps(res))
