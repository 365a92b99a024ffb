#!/usr/bin/env python
"""Where does the time of one MSDeformAttn MODULE call go?  Times the module's
forward+backward on an encoder-form and a decoder-form call and, separately, its pieces: the four
Linear layers, the elementwise prologue (softmax, location arithmetic), the fused op.
   python tools/module_bench.py [--dtype bf16|f32] [--iters 10]"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200.functions import MSDeformAttnFunction, MSDeformAttnFusedFunction  # noqa: E402
from kinet_b200.modules import MSDeformAttn  # noqa: E402
from kinet_b200.workloads import encoder_reference_points, resolve  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(2000000)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    amp = args.dtype == "bf16"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for tag in ("C2_enc", "C2_dec"):
        s = resolve(tag)
        N, S, Lq, L = s["N"], s["S"], s["Lq"], s["L"]
        mod = MSDeformAttn(256, L, 8, 4).to(dev)
        torch.manual_seed(0)
        with torch.no_grad():
            mod.sampling_offsets.weight.normal_(0, 0.01)
            mod.attention_weights.weight.normal_(0, 0.05)
        src = torch.randn(N, S, 256, device=dev, requires_grad=True)
        query = src if s["encoder"] else torch.randn(N, Lq, 256, device=dev, requires_grad=True)
        if s["encoder"]:
            ref = encoder_reference_points(s["levels"], N, dev).contiguous()
        else:
            ref = torch.rand(N, Lq, 1, 2, device=dev).expand(-1, -1, L, -1).contiguous()
        shapes = torch.as_tensor(s["levels"], dtype=torch.int64, device=dev)
        gout = torch.randn(N, Lq, 256, device=dev)

        def full():
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                out = mod(query, ref, src, shapes)
            out.backward(gout.to(out.dtype))

        # pieces
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            value = mod.value_proj(src).view(N, S, 8, 32).detach().requires_grad_(True)
            off = mod.sampling_offsets(query).view(N, Lq, 8, L, 4, 2).detach().requires_grad_(True)
            logit = mod.attention_weights(query).view(N, Lq, 8, L * 4).detach().requires_grad_(True)

        def prologue():
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                a = F.softmax(logit, -1).view(N, Lq, 8, L, 4)
                loc = ref[:, :, None, :, None, :] + off / shapes[None, None, None, :, None, :]
            return loc, a

        loc0, a0 = prologue()
        loc0 = loc0.detach().requires_grad_(True)
        a0 = a0.detach().requires_grad_(True)
        gl, ga = torch.randn_like(loc0), torch.randn_like(a0)

        def prologue_fb():
            loc, a = prologue()
            torch.autograd.backward([loc, a], [gl.to(loc.dtype), ga.to(a.dtype)])

        def op_fb():
            o = MSDeformAttnFunction.apply(value, shapes, loc0, a0, 64)
            o.backward(gout.to(o.dtype))

        def fused_fb():
            o = MSDeformAttnFusedFunction.apply(value, shapes, ref, off, logit, None, shapes, 64)
            o.backward(gout.to(o.dtype))

        def fused_f():
            with torch.no_grad():
                MSDeformAttnFusedFunction.apply(value, shapes, ref, off, logit, None, shapes, 64)

        def op_f():
            with torch.no_grad():
                MSDeformAttnFunction.apply(value, shapes, loc0, a0, 64)

        def linears_fb():
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                v = mod.value_proj(src)
                o = mod.sampling_offsets(query)
                w = mod.attention_weights(query)
                y = mod.output_proj(gout.to(v.dtype))
            torch.autograd.backward([v, o, w, y], [torch.ones_like(v), torch.ones_like(o), torch.ones_like(w), torch.ones_like(y)])

        mod.fused_prologue = False
        t_full_unfused = timeit(full, args.iters, flush)
        mod.fused_prologue = True
        t_full = timeit(full, args.iters, flush)
        print("        module fwd+bwd: fused prologue %8.1f us, PyTorch prologue %8.1f us" % (t_full, t_full_unfused), flush=True)
        t_pro = timeit(prologue_fb, args.iters, flush)
        t_op = timeit(op_fb, args.iters, flush)
        t_lin = timeit(linears_fb, args.iters, flush)
        t_fused = timeit(fused_fb, args.iters, flush)
        print("        fused op fwd+bwd %8.1f us (fwd %8.1f) | plain op fwd %8.1f | offsets dtype %s" % (
            t_fused, timeit(fused_f, args.iters, flush), timeit(op_f, args.iters, flush), off.dtype), flush=True)
        print("%-7s %-4s module fwd+bwd %8.1f us | op %8.1f  prologue(softmax+loc) %8.1f  4 linears %8.1f | loc dtype %s attn dtype %s value dtype %s"
              % (tag, args.dtype, t_full, t_op, t_pro, t_lin, loc0.dtype, a0.dtype, value.dtype), flush=True)


if __name__ == "__main__":
    main()
