"""Counters of the dense-window backward after one call: items claimed / done / left, flags of tiles and coarse chunks."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kinet_b200 import _lib
from kinet_b200.functions import ms_deform_attn_func as F
from kinet_b200.workloads import make_inputs, resolve

tag = sys.argv[1] if len(sys.argv) > 1 else "C2_enc"
regime = sys.argv[2] if len(sys.argv) > 2 else "structured"
s = resolve(tag)
dev = torch.device("cuda:0")
dt = torch.bfloat16
value, shapes, loc, attn = make_inputs(s, dev, dt, torch.float32, regime)
go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(dt)
lib = _lib.load()
aux, loc2, attn2, shapes2, lsi = F._prep(value, shapes, None, loc, attn)
dims = F._dims(value, loc2)
gv, gl, ga = torch.empty_like(value), torch.empty_like(loc2), torch.empty_like(attn2)
vcode = _lib.dtype_code(value.dtype)
ws_bytes = int(lib.msda_b200_backward_workspace_bytes(vcode, ctypes.byref(dims)))
ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=dev)
rc = lib.msda_b200_backward(F._ptr(value), vcode, F._ptr(shapes2), F._ptr(lsi), F._ptr(loc2), F._ptr(attn2), _lib.dtype_code(aux),
                            F._ptr(go), F._ptr(gv), F._ptr(gl), F._ptr(ga), F._ptr(ws), ws_bytes, ctypes.byref(dims), _lib.raw_stream(dev))
torch.cuda.synchronize()
nval = value.numel()
off = ((nval * 4 + 255) & ~255) // 4
cnt = ws[off:off + 4].view(torch.int32).tolist()
flags = ws[off + 64:].view(torch.uint8)
NM = s["N"] * s["M"]
nz = torch.nonzero(flags).flatten()
print(tag, regime, "rc", rc, "claims dense/rows/done/left", cnt, "flags set", int(flags.sum()), "last set flag", int(nz.max()) if len(nz) else None, "NM", NM)
per = flags[: (len(flags) // NM) * NM].view(-1, NM).sum(1)
print("flags per tile/chunk index:", per[:260].tolist())
