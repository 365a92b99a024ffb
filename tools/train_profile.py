#!/usr/bin/env python
"""Kernel breakdown of the data-parallel training step bench.py measures (product API: DeformableTransformer at
dropout 0.1, bf16 autocast, 800x1333 pyramid, batch 2), one eager step under the torch profiler.
    python tools/train_profile.py [--top 40]"""
import argparse
import collections
import os
import sys

import torch
from torch import nn
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kinet_b200.modules import DeformableTransformer  # noqa: E402
from kinet_b200.training import FlatParams, reference_param_groups  # noqa: E402
from kinet_b200.workloads import PYRAMIDS  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--top", type=int, default=40)
ap.add_argument("--dropout", type=float, default=0.1)
ap.add_argument("--ops", action="store_true", help="aten ops by input shape (where the elementwise tail comes from)")
ap.add_argument("--stacks", default="", help="comma-separated aten op names: print the Python call sites of their launches")
args = ap.parse_args()
dev = torch.device("cuda", 0)
levels, d = PYRAMIDS["detr_800x1333"], 256
torch.manual_seed(0)
tr = DeformableTransformer(d, 8, 6, 6, 1024, args.dropout, "relu", False, len(levels), 4, 4).to(dev).train()
qe = nn.Embedding(300, 2 * d).to(dev)
net = nn.ModuleList([tr, qe])
fp = FlatParams(net, groups=reference_param_groups(net))
opt = torch.optim.AdamW(fp.optimizer_groups(), lr=2e-4, weight_decay=1e-4, fused=True)
srcs = [torch.randn(2, d, h, w, device=dev) for h, w in levels]
pos = [torch.randn(2, d, h, w, device=dev) * 0.1 for h, w in levels]
masks = [torch.zeros(2, h, w, dtype=torch.bool, device=dev) for h, w in levels]
target = torch.randn(2, 300, d, device=dev)


def step():
    fp.detach_grads()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        hs = tr(srcs, masks, pos, qe.weight, None)[0]
    (hs.float() - target).square().mean().backward()
    fp.gather_grads()
    fp.clip_grad_norm_(0.1)
    opt.step()
    fp.refresh()


for _ in range(3):
    step()
torch.cuda.synchronize()
if args.stacks:
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=True) as prof:
        step()
        torch.cuda.synchronize()
    want = set(args.stacks.split(","))
    rows = [e for e in prof.key_averages(group_by_input_shape=True, group_by_stack_n=12) if e.key in want]
    for e in sorted(rows, key=lambda e: -e.device_time_total)[:args.top]:
        print("%-14s x%-3d %8.1f us  %s" % (e.key, e.count, e.device_time_total, str(e.input_shapes)[:100]))
        for fr in e.stack[:12]:
            if "kinet_b200" in fr or "tools/" in fr:
                print("      ", fr[-110:])
    sys.exit(0)
if args.ops:
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        step()
        torch.cuda.synchronize()
    print(prof.key_averages(group_by_input_shape=True).table(sort_by="self_device_time_total", row_limit=args.top,
                                                             max_name_column_width=50, max_shapes_column_width=90))
    sys.exit(0)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0.0, 0])
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        agg[e.name][0] += e.device_time
        agg[e.name][1] += 1
total = sum(v[0] for v in agg.values())
print("training step (dropout %.2f): %.0f us of kernels, %d launches" % (args.dropout, total, sum(v[1] for v in agg.values())))
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:args.top]:
    print("%9.1f us %5.1f %%  x%-4d %s" % (us, 100 * us / total, n, name[:120]))
