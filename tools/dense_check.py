#!/usr/bin/env python
"""Development check of the dense-window (tcgen05) encoder backward against the per-row kernel of the same
library (MSDA_B200_DENSE=0) and, for small shapes, the fp64 oracle.  python tools/dense_check.py [--time]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200.functions import ms_deform_attn_backward  # noqa: E402
from kinet_b200.workloads import make_inputs, resolve  # noqa: E402


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def run(value, shapes, loc, attn, go, dense):
    os.environ["MSDA_B200_DENSE"] = "1" if dense else "0"
    out = ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
    torch.cuda.synchronize()
    return out


def case(name, s, dt, regime, timeit):
    dev = torch.device("cuda:0")
    value, shapes, loc, attn = make_inputs(s, dev, dt, torch.float32, regime)
    go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(dt)
    a = run(value, shapes, loc, attn, go, True)
    b = run(value, shapes, loc, attn, go, False)
    msg = "%-28s %-5s %-10s gv %.3e  gl %.3e  ga %.3e" % (name, str(dt).split(".")[1], regime, rel(a[0], b[0]), rel(a[1], b[1]), rel(a[2], b[2]))
    if timeit:
        for dense in (True, False):
            os.environ["MSDA_B200_DENSE"] = "1" if dense else "0"
            for _ in range(3):
                ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
            e1.record()
            torch.cuda.synchronize()
            msg += "  %s %.1f us" % ("dense" if dense else "rows", e0.elapsed_time(e1) * 100)
    print(msg, flush=True)


def fused_case(name, s, dt, ref_dim, masked):
    """Module-prologue entry point (raw offsets / logits / reference points): dense vs per-row kernels."""
    from kinet_b200.functions import MSDeformAttnFusedFunction
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    N, S, M, D, L, P, Lq = s["N"], s["S"], s["M"], s["D"], s["L"], s["P"], s["Lq"]
    value = torch.randn(N, S, M, D, generator=g, device=dev).to(dt)
    shapes = torch.as_tensor(s["levels"], dtype=torch.int64, device=dev)
    from kinet_b200.workloads import encoder_reference_points
    ref = encoder_reference_points(s["levels"], N, dev).contiguous()
    if ref_dim == 4:
        wh = torch.rand(N, Lq, L, 2, generator=g, device=dev) * 0.1 + 0.02
        ref = torch.cat([ref, wh], -1).contiguous()
    off = (torch.randn(N, Lq, M, L, P, 2, generator=g, device=dev) * 2.0)
    logit = torch.randn(N, Lq, M, L * P, generator=g, device=dev)
    mask = (torch.rand(N, Lq, generator=g, device=dev) < 0.1) if masked else None
    go = torch.randn(N, Lq, M * D, generator=g, device=dev).to(dt)
    outs = []
    for dense in (True, False):
        os.environ["MSDA_B200_DENSE"] = "1" if dense else "0"
        v = value.clone().requires_grad_(True)
        r = ref.clone().requires_grad_(True)
        o = off.clone().requires_grad_(True)
        lg = logit.clone().requires_grad_(True)
        out = MSDeformAttnFusedFunction.apply(v, shapes, r, o, lg, mask, shapes, 64)
        out.backward(go)
        torch.cuda.synchronize()
        outs.append((v.grad, o.grad, lg.grad, r.grad))
    a, b = outs
    print("%-18s %-8s ref%d mask=%d  gv %.3e  goff %.3e  glogit %.3e  gref %.3e" % (
        name, str(dt).split(".")[1], ref_dim, int(masked), rel(a[0], b[0]), rel(a[1], b[1]), rel(a[2], b[2]), rel(a[3], b[3])), flush=True)


def main():
    timeit = "--time" in sys.argv
    small = dict(levels=[(20, 31), (10, 16), (5, 8), (3, 4)], N=2, M=2, D=32, P=4, encoder=True)
    small["L"] = 4
    small["S"] = sum(h * w for h, w in small["levels"])
    small["Lq"] = small["S"]
    odd = dict(levels=[(37, 53), (19, 27), (1, 9), (7, 1)], N=1, M=3, D=32, P=4, encoder=True)
    odd["L"] = 4
    odd["S"] = sum(h * w for h, w in odd["levels"])
    odd["Lq"] = odd["S"]
    d64 = dict(levels=[(24, 40), (12, 20)], N=1, M=2, D=64, P=3, encoder=True)
    d64["L"] = 2
    d64["S"] = sum(h * w for h, w in d64["levels"])
    d64["Lq"] = d64["S"]
    for name, s in (("small", small), ("odd", odd), ("d64", d64)):
        for dt in (torch.bfloat16, torch.float16):
            for regime in ("structured", "uniform"):
                case(name, s, dt, regime, False)
    for name, s in (("small", small), ("odd", odd), ("d64", d64)):
        for dt in (torch.bfloat16, torch.float16):
            for ref_dim, masked in ((2, False), (2, True), (4, True)):
                fused_case(name, s, dt, ref_dim, masked)
    fused_case("C2_enc", resolve("C2_enc"), torch.bfloat16, 2, False)
    for mc in [int(x) for x in os.environ.get("SWEEP_MAXCH", "5").split(",")]:
        os.environ["MSDA_B200_DENSE_MAXCH"] = str(mc)
        print("max level chunks", mc)
        for tag in ("C2_enc", "C5_enc"):
            for regime in ("structured", "uniform"):
                case(tag, resolve(tag), torch.bfloat16, regime, timeit)


if __name__ == "__main__":
    main()
