#!/usr/bin/env python
"""Where a tracking frame's GPU time goes (tools/frames_bench.py's workload, eager, torch profiler):
prints the kernels of one frame forward grouped by name, with their share of the summed kernel time.

    python tools/frame_profile.py [--train]     # --train: the training step of tools/train_bench.py instead
"""
import argparse
import collections
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _harness import FusedStack, Stack  # noqa: E402
from kinet_b200.workloads import PYRAMIDS  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--train", action="store_true")
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--layers", choices=("fused", "torch"), default="fused",
                    help="fused: kinet_b200.modules encoder/decoder classes; torch: the same layers as plain PyTorch ops")
    args = ap.parse_args()
    make_stack = FusedStack if args.layers == "fused" else Stack
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    if args.train:
        levels, d, nq, N = PYRAMIDS["detr_800x1333"], 256, 300, 2
    else:
        levels, d, nq, N = PYRAMIDS["mot17_750x1333"], 288, 560, 1
    S = sum(h * w for h, w in levels)
    model = make_stack(levels, n_queries=nq, d=d).to(dev)
    src = torch.randn(N, S, d, device=dev)
    pos = torch.randn(N, S, d, device=dev) * 0.1

    def run():
        if args.train:
            model.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out, _ = model(src, pos)
            out.float().square().mean().backward()
        else:
            with torch.no_grad():
                model(src, pos)

    if not args.train:      # one autocast region for the whole sequence (weight casts cached), as frames_bench.py
        torch.autocast("cuda", dtype=torch.bfloat16).__enter__()

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        run()
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0.0, 0])
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            agg[e.name][0] += e.device_time
            agg[e.name][1] += 1
    total = sum(v[0] for v in agg.values())
    print(f"{'train step' if args.train else 'frame forward'}: {total:.0f} us of kernels, "
          f"{sum(v[1] for v in agg.values())} launches")
    for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:args.top]:
        print(f"{us:9.1f} us {100 * us / total:5.1f} %  x{n:<4d} {name[:110]}")


if __name__ == "__main__":
    main()
