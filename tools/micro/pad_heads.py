"""D = 36 run as D = 64 with zero-padded heads: value_proj + MSDeformAttn forward + output_proj at the C3 shapes."""
import os, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kinet_b200.functions import ms_deform_attn_forward
from kinet_b200.workloads import make_inputs, resolve

dev = torch.device("cuda:0")
for tag in ("C3_enc", "C3_dec"):
    s = resolve(tag)
    N, S, M, D, Lq = s["N"], s["S"], s["M"], s["D"], s["Lq"]
    value, shapes, loc, attn = make_inputs(s, dev, torch.bfloat16, torch.float32, "structured")
    C = M * D
    x = torch.randn(N, S, C, device=dev, dtype=torch.bfloat16)
    wv = torch.randn(C, C, device=dev, dtype=torch.bfloat16) * 0.05
    bv = torch.randn(C, device=dev, dtype=torch.bfloat16) * 0.05
    wo = torch.randn(C, C, device=dev, dtype=torch.bfloat16) * 0.05
    bo = torch.randn(C, device=dev, dtype=torch.bfloat16) * 0.05
    DP = 64
    wv_p = torch.zeros(M, DP, C, device=dev, dtype=torch.bfloat16); wv_p[:, :D] = wv.view(M, D, C); wv_p = wv_p.view(M * DP, C)
    bv_p = torch.zeros(M, DP, device=dev, dtype=torch.bfloat16); bv_p[:, :D] = bv.view(M, D); bv_p = bv_p.view(-1)
    wo_p = torch.zeros(C, M, DP, device=dev, dtype=torch.bfloat16); wo_p[:, :, :D] = wo.view(C, M, D); wo_p = wo_p.view(C, M * DP)

    def plain():
        v = F.linear(x, wv, bv).view(N, S, M, D)
        o = ms_deform_attn_forward(v, shapes, loc, attn, 64)
        return F.linear(o, wo, bo)

    def padded():
        v = F.linear(x, wv_p, bv_p).view(N, S, M, DP)
        o = ms_deform_attn_forward(v, shapes, loc, attn, 64)
        return F.linear(o, wo_p, bo)

    a, b = plain(), padded()
    print(tag, "max abs diff", float((a.float() - b.float()).abs().max()), "of", float(a.float().abs().max()))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, fn in (("plain D=36", plain), ("padded D=64", padded)):
        g = torch.cuda.CUDAGraph()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            keep = fn()
        ts = []
        for _ in range(15):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        print("  %-12s %.1f us (value_proj + sampling + output_proj, median of 15, L2 flushed)" % (name, ts[7]))
