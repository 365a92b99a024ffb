#!/usr/bin/env python
"""Host-side time per phase of kinet_b200.tracking.SequenceTracker.run (perf_counter, no added syncs) next to the
device-timed frame: tells whether the tracking loop is bound by the device or by the host thread."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn  # noqa: E402

from kinet_b200.modules import DeformableTransformer  # noqa: E402
from kinet_b200.tracking import DetectionHeads, SequenceTracker  # noqa: E402
from kinet_b200.workloads import PYRAMIDS  # noqa: E402


def main(frames=200):
    dev = torch.device("cuda:0")
    levels, d, nq = PYRAMIDS["mot17_750x1333"], 288, 500
    torch.manual_seed(1000)
    tr = DeformableTransformer(d, 8, 6, 6, 1024, 0.1, "relu", True, len(levels), 4, 4).to(dev).eval()
    heads = DetectionHeads(d, 1).to(dev).eval()
    with torch.no_grad():
        heads.class_embed.bias.zero_()
        nn.init.normal_(heads.class_embed.weight, std=0.5)
        nn.init.normal_(heads.bbox_embed.layers[-1].weight, std=0.02)
    qe = torch.randn(nq, 2 * d, device=dev)
    g = torch.Generator(device=dev).manual_seed(7)
    scene = [torch.randn(1, d, h, w, device=dev, generator=g) for h, w in levels]
    pos = [torch.randn(1, d, h, w, device=dev, generator=g) * 0.1 for h, w in levels]
    masks = [torch.zeros(1, h, w, dtype=torch.bool, device=dev) for h, w in levels]
    feeds = [[s_ + 0.3 * torch.randn(s_.shape, device=dev, generator=g) for s_ in scene] for _ in range(8)]
    tracker = SequenceTracker(tr, heads, qe, img_size=(1080, 1920), use_graphs=True, bucket=32, autocast_dtype=torch.bfloat16)
    tracker.det_thresh = tracker.trk_thresh = 2.0
    tracker.step(feeds[0], masks, pos)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        hs, _, init_ref, inter_ref, _, _ = tr(feeds[0], masks, pos, qe, None)
        sc = heads(hs[-1], inter_ref[-2])[0].sigmoid().max(-1)[0].flatten()
    tracker.det_thresh = tracker.trk_thresh = float(torch.quantile(sc.float(), 0.92))
    tracker.reset()
    for i in range(24):
        tracker.step(feeds[i & 7], masks, pos)
    torch.cuda.synchronize()
    acc = {}
    for name in ("_encode_graph", "_decode_graph", "_heads_and_pack", "_update_tracks", "_targets"):
        fn = getattr(tracker, name)

        def wrap(*a, _fn=fn, _n=name, **k):
            t = time.perf_counter()
            r = _fn(*a, **k)
            acc[_n] = acc.get(_n, 0.0) + time.perf_counter() - t
            return r
        setattr(tracker, name, wrap)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in tracker.run((feeds[i & 7], masks, pos) for i in range(frames)):
        pass
    b.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    print("frames %d: device-timed %.3f ms/frame, wall %.3f ms/frame" % (frames, a.elapsed_time(b) / frames, wall / frames * 1e3))
    for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
        print("  host %-18s %.3f ms/frame" % (k, v / frames * 1e3))
    print("  host total of the above %.3f ms/frame" % (sum(acc.values()) / frames * 1e3))


if __name__ == "__main__":
    main()
