"""Per-frame CUDA graphs for tracking inference with a VARIABLE number of queries.

In the reference's tracker every frame runs the deformable transformer once with ``num_queries`` object queries
plus one track query per live track (src/trackformer/models/tracker.py:286-309, 536-542;
deformable_transformer.py:133-257), so the decoder's query count changes from frame to frame while the pyramid
does not.  ``FrameGraphs`` records

  * ONE graph of the encoder (fixed shapes), and
  * one graph of the decoder per BUCKET of query counts: the queries are padded up to the next multiple of
    ``bucket`` and the padding rows are switched off with the decoder's own ``query_attn_mask``
    (key_padding_mask of the self-attention, zeroed sampling weights in MSDeformAttn,
    deformable_transformer.py:371, 379-381) -- the mechanism the reference already uses for padded track
    queries in a batch, so real queries see exactly the keys they would see unpadded,

and replays them: no per-kernel launch cost, no device->host sync, nothing allocated after the first frame of a
bucket.  Inputs are copied into the graphs' static buffers; outputs are views of static buffers (valid until
the next call).  Weights are read from where they were at recording time: call ``reset()`` after changing
them.  Inference only (``torch.no_grad``); ``encoder`` / ``decoder`` are kinet_b200.modules'
DeformableTransformerEncoder / DeformableTransformerDecoder.
"""
import torch


class FrameGraphs:
    def __init__(self, encoder, decoder, spatial_shapes, bucket=32, autocast_dtype=torch.bfloat16, warmup=2, post=None):
        """post: optional ``post(hs, reference_points_out)`` recorded into every decoder graph behind the decoder (e.g.
        the detection heads); it sees the PADDED query rows; its result (static buffers, valid until the next call) is
        left in ``self.post_out`` by decode()."""
        if bucket <= 0:
            raise ValueError("bucket must be positive")
        self.encoder, self.decoder = encoder, decoder
        self.post, self.post_out = post, None
        self.shapes = spatial_shapes
        self.bucket = int(bucket)
        self.autocast_dtype = autocast_dtype
        self.warmup = int(warmup)
        self.reset()

    def reset(self):
        """Forget every recorded graph (after a weight update, or to release their memory)."""
        self._enc = None          # (graph, src_buf, pos_buf, valid_buf, mask_buf, memory stream)
        self._dec = {}            # padded query count -> (graph, static inputs, out, ref_out)
        self.replays = 0

    @property
    def recorded_buckets(self):
        return sorted(self._dec)

    def _amp(self):
        return torch.autocast("cuda", dtype=self.autocast_dtype, enabled=self.autocast_dtype is not None)

    def _record(self, fn):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):     # fills the per-tensor caches (shape check, level grids, weight copies)
                fn()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = fn()
        return graph, out

    def _encode(self, src, pos, valid_ratios, padding_mask):
        if self._enc is None or self._enc[1].shape != src.shape:
            bufs = (torch.empty_like(src), torch.empty_like(pos), torch.empty_like(valid_ratios),
                    None if padding_mask is None else torch.empty_like(padding_mask))
            for b, t in zip(bufs, (src, pos, valid_ratios, padding_mask)):
                if b is not None:
                    b.copy_(t)

            def fn():
                with torch.no_grad(), self._amp():
                    return self.encoder(bufs[0], self.shapes, bufs[2], bufs[1], bufs[3], return_stream=True)
            graph, mem = self._record(fn)
            self._enc = (graph,) + bufs + (mem,)
            self._dec = {}
        graph, src_b, pos_b, valid_b, mask_b, mem = self._enc
        src_b.copy_(src, non_blocking=True)
        pos_b.copy_(pos, non_blocking=True)
        valid_b.copy_(valid_ratios, non_blocking=True)
        if mask_b is not None:
            mask_b.copy_(padding_mask, non_blocking=True)
        graph.replay()
        return mem, valid_b, mask_b

    def encode(self, src, pos, valid_ratios, padding_mask=None):
        """Replay the encoder graph on a new frame.  The decoder graphs read its output buffer, so the call may be
        issued as soon as the previous frame's decode() has been ENQUEUED (same stream): a tracker can start frame
        t+1's encoder while the host is still reading frame t's results back."""
        self._enc_out = self._encode(src, pos, valid_ratios, padding_mask)

    def decode(self, tgt, reference_points, query_pos, query_attn_mask=None):
        """Replay the decoder graph of the bucket that fits the query count, on the memory of the last encode()."""
        return self._decode(self._enc_out, tgt, reference_points, query_pos, query_attn_mask)

    def __call__(self, src, pos, valid_ratios, tgt, reference_points, query_pos, padding_mask=None,
                 query_attn_mask=None):
        """src, pos (N, S, C); valid_ratios (N, L, 2); tgt, query_pos (N, Lq, C); reference_points (N, Lq, 2|4);
        padding_mask (N, S) bool or None (the choice is fixed at the first call); query_attn_mask (N, Lq) bool
        or None.  Returns (hs, reference_points_out, memory) like decoder(...) + encoder(...): hs is
        (N, Lq, C), or (layers, N, Lq, C) when the decoder returns intermediates."""
        return self._decode(self._encode(src, pos, valid_ratios, padding_mask), tgt, reference_points, query_pos,
                            query_attn_mask)

    def _decode(self, enc_out, tgt, reference_points, query_pos, query_attn_mask):
        N, Lq, C = tgt.shape
        mem, valid_b, mask_b = enc_out
        B = max(self.bucket, -(-Lq // self.bucket) * self.bucket)
        if B not in self._dec:
            dev = tgt.device
            tgt_b = torch.zeros(N, B, C, dtype=tgt.dtype, device=dev)
            qpos_b = torch.zeros(N, B, C, dtype=query_pos.dtype, device=dev)
            ref_b = torch.full((N, B, reference_points.shape[-1]), 0.5, dtype=reference_points.dtype, device=dev)
            qmask_b = torch.ones(N, B, dtype=torch.bool, device=dev)
            qmask_b[:, :Lq] = False if query_attn_mask is None else query_attn_mask
            tgt_b[:, :Lq], qpos_b[:, :Lq], ref_b[:, :Lq] = tgt, query_pos, reference_points

            def fn():
                with torch.no_grad(), self._amp():
                    hs_, ref_ = self.decoder(tgt_b, ref_b, mem, self.shapes, valid_b, qpos_b, mask_b, qmask_b)
                with torch.no_grad():
                    return hs_, ref_, (self.post(hs_, ref_) if self.post is not None else None)
            graph, (hs, ref_out, extra) = self._record(fn)
            self._dec[B] = (graph, tgt_b, qpos_b, ref_b, qmask_b, hs, ref_out, extra)
        graph, tgt_b, qpos_b, ref_b, qmask_b, hs, ref_out, extra = self._dec[B]
        tgt_b[:, :Lq].copy_(tgt, non_blocking=True)
        qpos_b[:, :Lq].copy_(query_pos, non_blocking=True)
        ref_b[:, :Lq].copy_(reference_points, non_blocking=True)
        qmask_b.fill_(True)
        if query_attn_mask is None:
            qmask_b[:, :Lq] = False
        else:
            qmask_b[:, :Lq].copy_(query_attn_mask, non_blocking=True)
        graph.replay()
        self.replays += 1
        self.post_out = extra
        return hs[..., :Lq, :], ref_out[..., :Lq, :], mem.x


class KineFrameGraphs:
    """Per-frame CUDA graphs for the `kine` detector (modules.kinematic_detector.KinematicDetectorTransformer around
    modules.kinematic_transformer.DualKinematicTransformer; reference detr.py:288-417, transformer.py:145-185).

    A frame brings T detection tokens and K track queries, both changing from frame to frame.  Tokens are padded up to
    a multiple of ``token_bucket`` and switched off with the model's own key-padding ``mask`` (what the reference's
    NestedTensor batching does); track queries are padded up to a multiple of ``track_bucket`` at the FRONT of the
    query list with rows that nothing reads back, flagged in ``bufs["kpad"]`` (1, Kp) -- a ``heads`` function that hands
    that flag to ``DualKinematicTransformer.forward_masked_queries`` (padded queries masked as keys of the decoder's
    self-attention) gets exactly the unpadded results for the real queries; one that ignores it lets the padding rows
    take part as keys.  One graph per (padded T, padded K); inputs are copied into static buffers, outputs are views of
    static buffers valid until the next call.  Inference only."""

    def __init__(self, transformer, heads, token_bucket=32, track_bucket=1, autocast_dtype=torch.bfloat16, warmup=2):
        """transformer: modules.kinematic_transformer.DualKinematicTransformer; heads: see __call__."""
        self.transformer, self.heads = transformer, heads
        self.token_bucket, self.track_bucket = int(token_bucket), int(track_bucket)
        self.autocast_dtype, self.warmup = autocast_dtype, int(warmup)
        self._graphs = {}
        self.replays = 0

    def reset(self):
        self._graphs = {}

    @property
    def recorded_buckets(self):
        return sorted(self._graphs)

    def _record(self, key, bufs, pre=None):
        Kp = key[1]

        def fn():
            with torch.no_grad(), torch.autocast("cuda", dtype=self.autocast_dtype,
                                                 enabled=self.autocast_dtype is not None):
                if pre is not None:
                    pre()
                return self.heads(self.transformer, bufs, Kp)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = fn()
        self._graphs[key] = (graph, bufs, out)

    def call_with_trails(self, src_det, src_meta, pos_det, pos_meta, trail_det, trail_meta):
        """The tracking loop's entry: the K track queries arrive as HOST arrays ``trail_det`` (K, Xd) / ``trail_meta``
        (K, Xm) (numpy float32; K may be 0) and everything that depends on the frame's counts happens inside the graph.
        The token count, the number of padding rows and the (front-padded) trails travel in ONE pinned block and one
        host -> device copy straight into the bucket's static buffer; the graph derives ``mask`` / ``kpad`` from the two
        counts and ``heads`` finds the raw trails in ``bufs["trail_det"]`` (Kp, Xd) / ``bufs["trail_meta"]`` (Kp, Xm) --
        projecting them to track-query embeddings is part of what it records.  Per frame the host enqueues the four
        token copies, that block and the replay: 6 launches instead of ~25.  Same return value as __call__."""
        T, C = src_det.shape[1], src_det.shape[2]
        graph, bufs, out, pad = self._block_call(src_det.device, T, (), trail_det, trail_meta, tokens_c=C)
        for name, t in (("src_det", src_det), ("src_meta", src_meta), ("pos_det", pos_det), ("pos_meta", pos_meta)):
            bufs[name][:, :T].copy_(t, non_blocking=True)
        graph.replay()
        self.replays += 1
        return out, pad

    def call_with_detections(self, device, dets, meta, trail_det, trail_meta):
        """A frame from its RAW detections: ``dets`` (T, Xb) boxes and ``meta`` (T, Xc) confidences [+ classes] as HOST
        arrays travel in the same pinned block as the counts and the trails; ``heads`` finds them in ``bufs["det_raw"]``
        (1, Tp, Xb) / ``bufs["meta_raw"]`` (1, Tp, Xc) next to ``mask`` and runs the model's backbones
        (modules.kinematic_backbone) as part of the recorded graph.  The host enqueues ONE copy and the replay."""
        graph, bufs, out, pad = self._block_call(device, dets.shape[0], (dets, meta), trail_det, trail_meta)
        graph.replay()
        self.replays += 1
        return out, pad

    def _block_call(self, dev, T, raw, trail_det, trail_meta, tokens_c=0):
        """Shared by the two entries above: finds / records the graph of the frame's bucket and sends the pinned block
        [T, padding track rows | raw detections, raw metadata (Tp rows each, zero-padded at the END) | trails (Kp rows each,
        zero-padded at the FRONT)].  -> (graph, bufs, out, number of padding track-query rows)."""
        K, Xd, Xm = trail_det.shape[0], trail_det.shape[1], trail_meta.shape[1]
        Tp = max(-(-T // self.token_bucket), 1) * self.token_bucket       # a frame without detections: one bucket, all masked
        Kp = -(-K // self.track_bucket) * self.track_bucket
        Xr = tuple(int(r.shape[1]) for r in raw)
        key = (Tp, Kp, Xd, Xm) + Xr
        n_raw = Tp * sum(Xr)
        n_block = 2 + n_raw + Kp * (Xd + Xm)
        if key not in self._graphs:
            block = torch.zeros(n_block, device=dev)
            o = 2 + n_raw
            bufs = dict(block=block, trail_det=block[o:o + Kp * Xd].view(Kp, Xd), trail_meta=block[o + Kp * Xd:].view(Kp, Xm))
            if raw:
                bufs["det_raw"] = block[2:2 + Tp * Xr[0]].view(1, Tp, Xr[0])
                bufs["meta_raw"] = block[2 + Tp * Xr[0]:2 + n_raw].view(1, Tp, Xr[1])
            else:
                for name in ("src_det", "src_meta", "pos_det", "pos_meta"):
                    bufs[name] = torch.zeros(1, Tp, tokens_c, device=dev)
            # (kept in bufs: the recorded graph reads them on every replay, so they must outlive this call)
            tok = bufs["token_index"] = torch.arange(Tp, device=dev, dtype=torch.float32).view(1, Tp)
            trk = bufs["track_index"] = torch.arange(Kp, device=dev, dtype=torch.float32).view(1, Kp)

            def pre():
                bufs["mask"] = tok >= block[0]           # tokens T .. Tp-1 are padding
                bufs["kpad"] = trk < block[1]            # the first Kp-K track-query rows are padding
            self._record(key, bufs, pre)
        graph, bufs, out = self._graphs[key]
        st = getattr(self, "_stage", None)
        if st is None or st[0].shape[0] < n_block:
            st = self._stage = [torch.zeros(max(4096, 2 * n_block)).pin_memory() for _ in range(2)] + [0]
        st[2] ^= 1                                      # two blocks in turn: the previous frame's copy may still be in flight
        host = st[st[2]]
        flat = host.numpy()
        flat[0], flat[1] = T, Kp - K
        o = 2
        for a, X in zip(raw, Xr):
            flat[o:o + T * X] = a.ravel()
            flat[o + T * X:o + Tp * X] = 0.0
            o += Tp * X
        if Kp:
            for a, X in ((trail_det, Xd), (trail_meta, Xm)):
                flat[o:o + (Kp - K) * X] = 0.0
                flat[o + (Kp - K) * X:o + Kp * X] = a.ravel()
                o += Kp * X
        bufs["block"].copy_(host[:n_block], non_blocking=True)
        return graph, bufs, out, Kp - K

    def __call__(self, src_det, src_meta, pos_det, pos_meta, tgt_det=None, tgt_meta=None):
        """src_* / pos_* (1, T, C) token embeddings; tgt_* (K, 1, C) projected track-query embeddings or None.
        ``heads(transformer, bufs, padded_K)`` is the frame's computation as the caller wants it recorded: it reads the
        static buffers ``bufs`` (src_det, src_meta, pos_det, pos_meta (1, Tp, C); mask (1, Tp) bool; tgt_det, tgt_meta
        (Kp, 1, C); kpad (1, Kp) bool = padding track-query rows), calls the transformer and returns whatever should come back (tools/kine_bench.py::make_heads: class
        logits, boxes and the detection embeddings of the last layer).  Returns (that result -- static buffers, valid
        until the next call --, number of padding track-query rows at the FRONT of the query dimension)."""
        T, C = src_det.shape[1], src_det.shape[2]
        K = 0 if tgt_det is None else tgt_det.shape[0]
        Tp = -(-T // self.token_bucket) * self.token_bucket
        Kp = -(-K // self.track_bucket) * self.track_bucket
        key = (Tp, Kp)
        dev = src_det.device
        if key not in self._graphs:
            bufs = dict(src_det=torch.zeros(1, Tp, C, device=dev), src_meta=torch.zeros(1, Tp, C, device=dev),
                        pos_det=torch.zeros(1, Tp, C, device=dev), pos_meta=torch.zeros(1, Tp, C, device=dev),
                        mask=torch.ones(1, Tp, dtype=torch.bool, device=dev),
                        kpad=torch.zeros(1, Kp, dtype=torch.bool, device=dev),
                        tgt_det=torch.zeros(Kp, 1, C, device=dev), tgt_meta=torch.zeros(Kp, 1, C, device=dev))

            self._record(key, bufs)
        graph, bufs, out = self._graphs[key]
        for name, t in (("src_det", src_det), ("src_meta", src_meta), ("pos_det", pos_det), ("pos_meta", pos_meta)):
            bufs[name][:, :T].copy_(t, non_blocking=True)
        bufs["mask"][:, :T] = False
        if Tp > T:
            bufs["mask"][:, T:] = True
        if K:
            bufs["tgt_det"][Kp - K:].copy_(tgt_det, non_blocking=True)
            bufs["tgt_meta"][Kp - K:].copy_(tgt_meta, non_blocking=True)
        if Kp:
            bufs["kpad"].fill_(False)
        if Kp > K:
            bufs["tgt_det"][:Kp - K].zero_()
            bufs["tgt_meta"][:Kp - K].zero_()
            bufs["kpad"][:, :Kp - K] = True
        graph.replay()
        self.replays += 1
        return out, Kp - K
