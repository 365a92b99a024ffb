#!/usr/bin/env python
"""Per-shape kernel timing (CUDA events, L2 flushed between iterations).  Development tool;
bench.py is the contract benchmark.   python tools/quick_bench.py [tags...] [--dtype bf16] [--iters 20]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200.functions import ms_deform_attn_backward, ms_deform_attn_forward  # noqa: E402
from kinet_b200.workloads import SHAPES, algorithmic_bytes, make_inputs, resolve  # noqa: E402


def time_op(fn, flush, iters, warmup=3):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()                       # > L2: evicts inputs
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(400000)           # let the host run ahead: events then bracket GPU time only
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tags", nargs="*", default=list(SHAPES))
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16", "f16"])
    ap.add_argument("--aux", default="f32", choices=["f32", "same"])
    ap.add_argument("--regime", default="structured")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-bwd", action="store_true")
    ap.add_argument("--warm", action="store_true", help="do not flush L2 between iterations")
    args = ap.parse_args()
    dt = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}[args.dtype]
    aux = dt if args.aux == "same" else torch.float32
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if not args.warm else torch.empty(1, device=dev)
    peak = 6550.7
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for tag in args.tags:
        s = resolve(tag)
        value, shapes, loc, attn = make_inputs(tag, dev, dt, aux, args.regime)
        go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(dt)
        bf, bb = algorithmic_bytes(s, value.element_size(), loc.element_size())
        med, best = time_op(lambda: ms_deform_attn_forward(value, shapes, loc, attn, 64), flush, args.iters)
        line = "%-7s %-4s %-10s fwd %8.1f us (best %8.1f)  %7.1f GB/s  %.3f of %.0f" % (
            tag, args.dtype, args.regime, med, best, bf / med / 1e3, bf / med / 1e3 / peak, peak)
        if not args.no_bwd:
            med, best = time_op(lambda: ms_deform_attn_backward(value, shapes, loc, attn, go, 64), flush, args.iters)
            line += " | bwd %8.1f us (best %8.1f)  %7.1f GB/s  %.3f" % (
                med, best, bb / med / 1e3, bb / med / 1e3 / peak)
        print(line, flush=True)


if __name__ == "__main__":
    main()
