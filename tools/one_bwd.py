#!/usr/bin/env python
"""A few encoder-form backward calls of one shape (profiling target).  python tools/one_bwd.py [tag] [dtype] [regime] [n]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200.functions import ms_deform_attn_backward, ms_deform_attn_forward  # noqa: E402
from kinet_b200.workloads import make_inputs, resolve  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "C2_enc"
dt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[sys.argv[2] if len(sys.argv) > 2 else "bf16"]
regime = sys.argv[3] if len(sys.argv) > 3 else "structured"
n = int(sys.argv[4]) if len(sys.argv) > 4 else 3
s = resolve(tag)
dev = torch.device("cuda:0")
value, shapes, loc, attn = make_inputs(s, dev, dt, torch.float32, regime)
go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(dt)
for _ in range(n):
    ms_deform_attn_forward(value, shapes, loc, attn, 64)
    ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
torch.cuda.synchronize()
print("ok")
