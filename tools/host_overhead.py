#!/usr/bin/env python
"""Host-side cost of one op call (Python wrapper + ctypes + launch), measured with tiny inputs so the GPU
is never the limiter.  usage: python tools/host_overhead.py [--profile]"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200.functions import MSDeformAttnFunction, ms_deform_attn_backward, ms_deform_attn_forward  # noqa: E402
from kinet_b200.workloads import make_inputs  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    shape = dict(N=1, S=20 * 30 + 10 * 15, M=8, D=32, L=2, P=4, Lq=64, levels=[(20, 30), (10, 15)], encoder=False)
    value, shapes, loc, attn = make_inputs(shape, dev)
    go = torch.randn(1, 64, 256, device=dev)
    n = 2000

    def timeit(fn, label):
        for _ in range(50):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("%-44s %7.1f us per call (host)" % (label, (t1 - t0) / n * 1e6))

    timeit(lambda: ms_deform_attn_forward(value, shapes, loc, attn, 64), "ms_deform_attn_forward")
    timeit(lambda: ms_deform_attn_backward(value, shapes, loc, attn, go, 64), "ms_deform_attn_backward")
    v = value.clone().requires_grad_(True)
    l = loc.clone().requires_grad_(True)
    a = attn.clone().requires_grad_(True)

    def fb():
        MSDeformAttnFunction.apply(v, shapes, l, a, 64).backward(go)
    timeit(fb, "MSDeformAttnFunction.apply + backward")
    timeit(lambda: torch.empty(1, 64, 256, device=dev), "torch.empty (for scale)")
    timeit(lambda: value + 1, "value + 1 (one eager torch op, for scale)")
    if "--profile" in sys.argv:
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(2000):
            ms_deform_attn_forward(value, shapes, loc, attn, 64)
        pr.disable()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(18)


if __name__ == "__main__":
    main()
