#!/usr/bin/env python
"""Tracking-configuration throughput of the transformer part (BASELINE.json configs[2], SURVEY.md C3 shapes):
per frame one 6-layer MSDeformAttn encoder over the MOT17 pyramid (S = 20 906, d = 288 -> D = 36) and one
6-layer decoder with 500 object + 60 track queries, bf16 autocast, forward only, recorded into ONE CUDA graph
per GPU and replayed frame after frame.  Sequences are independent, so N GPUs run N replicas with no
communication (engine.py:290-296 assigns sequences round-robin); frames/s is the sum over GPUs.
The backbone, heads and the tracker's Python are the reference's and are not part of this number.

    python tools/frames_bench.py [--frames 200]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/frames_bench.py
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _harness import FusedStack, Stack  # noqa: E402  (encoder/decoder stack around kinet_b200.MSDeformAttn)
from kinet_b200.workloads import PYRAMIDS  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=200)
    ap.add_argument("--eager", action="store_true", help="no CUDA graph (launch-bound reference point)")
    ap.add_argument("--stream", choices=("f32", "bf16"), default="f32",
                    help="dtype of the residual stream between layers: f32 = what the reference's modules do under "
                         "autocast (default); bf16 = feed bf16 activations, every layer then reads and writes half the bytes")
    ap.add_argument("--varying-queries", action="store_true",
                    help="500 object queries + a number of track queries that changes every frame (20..60), through "
                         "kinet_b200.frame_graph.FrameGraphs (one decoder graph per bucket of 32 queries)")
    ap.add_argument("--layers", choices=("fused", "torch"), default="fused",
                    help="fused: kinet_b200.modules encoder/decoder classes; torch: the same layers as plain PyTorch ops")
    args = ap.parse_args()
    make_stack = FusedStack if args.layers == "fused" else Stack
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    levels = PYRAMIDS["mot17_750x1333"]
    S = sum(h * w for h, w in levels)
    torch.manual_seed(rank)
    model = make_stack(levels, n_queries=560, d=288).to(dev).eval()
    src = torch.randn(1, S, 288, device=dev)
    pos = torch.randn(1, S, 288, device=dev) * 0.1
    frames = [torch.randn(1, S, 288, device=dev) for _ in range(4)]   # stand-ins for the backbone's output
    if args.stream == "bf16":
        assert args.layers == "fused"
        src, pos, frames = src.bfloat16(), pos.bfloat16(), [f.bfloat16() for f in frames]
        model.query_embed.to(torch.bfloat16)        # tgt / query_pos start the decoder's stream

    # ONE autocast region around warm-up, capture and replay: autocast's cache of bf16 weight copies lives as
    # long as the outermost region, so the per-frame graph holds no weight casts (a tracker loop would wrap
    # the whole sequence the same way)
    amp = torch.autocast("cuda", dtype=torch.bfloat16)
    amp.__enter__()

    def frame():
        with torch.no_grad():
            return model(src, pos)

    if args.varying_queries:
        from kinet_b200.frame_graph import FrameGraphs
        assert args.layers == "fused"
        shapes = torch.as_tensor(levels, dtype=torch.long, device=dev)
        valid = torch.ones(1, len(levels), 2, device=dev)
        fg = FrameGraphs(model.encoder, model.decoder, shapes, bucket=32)
        counts = [500 + 20 + (7 * i) % 41 for i in range(16)]                  # 520 .. 560 queries
        qs = [(torch.randn(1, n, 288, device=dev), torch.rand(1, n, 2, device=dev) * 0.8 + 0.1,
               torch.randn(1, n, 288, device=dev) * 0.1) for n in counts]
        amp.__exit__(None, None, None)           # FrameGraphs manages autocast itself

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(0 if args.varying_queries else 3):
            frame()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    if args.varying_queries:
        step = None
    elif args.eager:
        step = frame
    else:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            frame()
        step = graph.replay

    def run(n):
        for i in range(n):
            if args.varying_queries:
                tgt, ref, qpos = qs[i & 15]
                fg(frames[i & 3], pos, valid, tgt, ref, qpos)
                continue
            src.copy_(frames[i & 3], non_blocking=True)     # a new frame lands in the static input buffer
            step()

    run(10)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    run(args.frames)
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b)], device=dev)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"metric": "tracking_transformer_frames_per_s", "value": world * args.frames / ms * 1e3,
                          "n_gpus": world, "ms_per_frame_per_gpu": ms / args.frames, "frames_per_gpu": args.frames,
                          "mode": "frame_graphs_varying_queries" if args.varying_queries else ("eager" if args.eager else "cuda_graph"), "layers": args.layers, "residual_stream": args.stream,
                          "config": "MOT17 750x1333 pyramid, S=20906, d=288 (D=36), 6 enc + 6 dec layers, "
                                    "560 queries, bf16 autocast, forward; replicas, no communication"}))
    if not args.varying_queries:
        amp.__exit__(None, None, None)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
