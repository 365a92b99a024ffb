#!/usr/bin/env python
"""Per-phase cycle shares of the dense-window backward (needs a build with MSDA_NVCC_EXTRA=-DMSDA_DENSE_TIMING)."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kinet_b200 import _lib  # noqa: E402
from kinet_b200.functions import ms_deform_attn_backward  # noqa: E402
from kinet_b200.workloads import make_inputs, resolve  # noqa: E402

NAMES = ["item end->claim", "claim+decode+softmax", "bbox+windows", "G/Gt + level setup", "V load issue", "scatter W",
         "V store", "fence+barrier", "MMA issue", "wait U", "stage+picks", "wait GV", "GV epilogue", "clear W",
         "level finalize", "tail"]
tag = sys.argv[1] if len(sys.argv) > 1 else "C2_enc"
s = resolve(tag)
dev = torch.device("cuda:0")
value, shapes, loc, attn = make_inputs(s, dev, torch.bfloat16, torch.float32, "structured")
go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(torch.bfloat16)
lib = _lib.load()
buf = (ctypes.c_ulonglong * 16)()
for _ in range(3):
    ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
torch.cuda.synchronize()
lib.msda_b200_debug_dense_phase_cycles(buf)
ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
torch.cuda.synchronize()
lib.msda_b200_debug_dense_phase_cycles(buf)
tot = sum(buf)
print("total CTA-cycles %d (per CTA %.0f)" % (tot, tot / 296.0))
for n, v in zip(NAMES, buf):
    print("%-24s %12d  %5.1f%%" % (n, v, 100.0 * v / max(tot, 1)))
