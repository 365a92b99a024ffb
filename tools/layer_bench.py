#!/usr/bin/env python
"""CUDA-event timings of the layer glue kernels (csrc/layer_ops.cu) at the shapes of BASELINE.json configs[1]
(N=2, S=22 223, C=256) and configs[2] (N=1, S=20 906, C=288), with their algorithmic bytes against the
measured HBM peak, next to the PyTorch ops they replace.  Each candidate is replayed from a CUDA graph (device
time only), L2 flushed between iterations.

    python tools/layer_bench.py [--iters 20]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kinet_b200.functions.layer_ops import AddLayerNormFunction, column_sum  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6550.0


def timeit(fn, iters, flush):
    """Median device time of fn's kernels: fn is recorded into a CUDA graph once (no per-launch host cost in
    the number), every timed replay starts with a cold L2."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        keep = fn()                                   # noqa: F841  (outputs stay alive with the graph)
    ts = []
    for _ in range(iters):
        flush.zero_()                # write more than L2 ...
        flush[: 192 << 20].max()     # ... then read it back, so the candidate does not pay for evicting dirty lines
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); graph.replay(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    pk = peak()
    print(f"HBM peak (measured) {pk:.0f} GB/s; times are medians of {args.iters}, L2 flushed")
    for tag, rows, C in (("C2 N=2 S=22223 C=256", 44446, 256), ("C3 N=1 S=20906 C=288", 20906, 288)):
        g = torch.Generator(device=dev).manual_seed(0)
        x = torch.randn(rows, C, device=dev, generator=g)
        r = torch.randn(rows, C, device=dev, generator=g).bfloat16()
        pos = torch.randn(rows, C, device=dev, generator=g)
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        bf = torch.bfloat16
        for name, use_pos in (("norm(x+r) -> y f32, y bf16", False), ("... + (y+pos) bf16", True)):
            nbytes = rows * C * (4 + 2 + 4 + 2 + ((4 + 2) if use_pos else 0))
            us = timeit(lambda: AddLayerNormFunction.apply(x, r, pos if use_pos else None, gamma, beta, 1e-5, bf, True),
                        args.iters, flush)

            def torch_ops():
                y = F.layer_norm(x + r, (C,), gamma, beta, 1e-5)
                out = [y, y.to(bf)]
                if use_pos:
                    out.append((y + pos).to(bf))
                return out
            us0 = timeit(torch_ops, args.iters, flush)
            print(f"{tag}  fwd {name:32s} {us:7.1f} us  {nbytes / us / 1e3:6.0f} GB/s  {nbytes / us / 1e3 / pk:5.2f} of HBM   "
                  f"(PyTorch ops {us0:7.1f} us)")
        # backward: dy f32 + dy_lp bf16 + dypos bf16 + x + r -> dx f32, dr bf16, dgamma, dbeta.  Autograd cannot
        # replay a backward alone from a graph, so forward+backward is recorded and the forward's time subtracted.
        xg, rg = x.clone().requires_grad_(True), r.clone().requires_grad_(True)
        gg, bg = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        dy, dlp = torch.randn(rows, C, device=dev, generator=g), torch.randn(rows, C, device=dev, generator=g).to(bf)
        nbytes = rows * C * (4 + 2 + 2 + 4 + 2 + 4 + 2)

        def ours(backward):
            outs = AddLayerNormFunction.apply(xg, rg, pos, gg, bg, 1e-5, bf, True)
            return torch.autograd.grad(outs, (xg, rg, gg, bg), (dy, dlp, dlp)) if backward else outs

        def stock(backward):
            y0 = F.layer_norm(xg + rg, (C,), gg, bg, 1e-5)
            outs = (y0, y0.to(bf), (y0 + pos).to(bf))
            return torch.autograd.grad(outs, (xg, rg, gg, bg), (dy, dlp, dlp)) if backward else outs
        us = timeit(lambda: ours(True), args.iters, flush) - timeit(lambda: ours(False), args.iters, flush)
        us0 = timeit(lambda: stock(True), args.iters, flush) - timeit(lambda: stock(False), args.iters, flush)
        print(f"{tag}  bwd {'all three gradients in':32s} {us:7.1f} us  {nbytes / us / 1e3:6.0f} GB/s  "
              f"{nbytes / us / 1e3 / pk:5.2f} of HBM   (PyTorch ops {us0:7.1f} us)")
        for width in (C, 1024):
            a = torch.randn(rows, width, device=dev, generator=g).bfloat16()
            nbytes = rows * width * 2
            us = timeit(lambda: column_sum(a, torch.float32), args.iters, flush)
            us0 = timeit(lambda: a.sum(0), args.iters, flush)
            print(f"{tag}  column sum bf16 [{rows} x {width:4d}]{'':10s} {us:7.1f} us  {nbytes / us / 1e3:6.0f} GB/s  "
                  f"{nbytes / us / 1e3 / pk:5.2f} of HBM   (torch sum(0) {us0:7.1f} us)")


if __name__ == "__main__":
    main()
