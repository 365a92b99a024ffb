#!/usr/bin/env python
"""SASS evidence for profiles/: for the dominant kernels of libmsda_b200.so, the opcode counts that matter and the
first occurrences of the instructions DESIGN.md refers to (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,
LDGSTS = cp.async, REDG.E.ADD.F32x4 = red.global.add.v4.f32, RED...BF16x2, FFMA2, CCTL.E.PF2 = prefetch.global.L2,
SYNCS = mbarrier).    python tools/sass_excerpts.py > profiles/r2_sass_excerpts.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "kinet_b200", "libmsda_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
KERNELS = [("dense-window backward, bf16 value / f32 aux, D=32, plain", "msda_backward_dense_kernelI13__nv_bfloat16fLi32ELb0E"),
           ("dense-window backward, bf16 value / f32 aux, D=32, fused prologue", "msda_backward_dense_kernelI13__nv_bfloat16fLi32ELb1E"),
           ("per-row backward, bf16 value / f32 aux, VEC=4 GS=8 NW=8", "msda_backward_kernelI13__nv_bfloat16fLi4ELi8ELi8ELb0E"),
           ("forward, bf16 value / f32 aux, VEC=8 GS=4 NW=8", "msda_forward_kernelI13__nv_bfloat16fLi8ELi4ELi8ELb0E"),
           ("add + LayerNorm forward", "add_layernorm_fwd_kernel"),
           ("add + LayerNorm backward", "add_layernorm_bwd_kernel")]
SHOW = ["UTCHMMA", "UTCBAR", "LDTM", "UTCATOMSWS", "LDGSTS", "REDG.E.ADD.F32x4", "RED.E.ADD", "REDG.E.ADD.BF16x2", "FFMA2", "CCTL", "SYNCS",
        "FENCE.VIEW.ASYNC", "ATOMG", "LDG.E.128", "STS.128", "LDS.128", "REDUX", "ACQBULK", "PREEXIT"]
for title, sub in KERNELS:
    blocks = [b for b in txt.split("Function : ")[1:] if sub in b.split("\n", 1)[0]]
    if not blocks:
        print("== %s: not found (%s)\n" % (title, sub))
        continue
    blk = blocks[0]
    name = blk.split("\n", 1)[0].strip()
    lines = []
    for ln in blk.split("\n"):
        m = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', ln)
        if m:
            lines.append((int(m.group(1), 16), re.sub(r'\s+', ' ', m.group(2)).strip()))
    c = collections.Counter()
    for a, t in lines:
        toks = t.split()
        op = toks[1] if toks[0].startswith('@') and len(toks) > 1 else toks[0]
        c[op] += 1
    print("== %s\n   %s\n   %d instructions" % (title, demangle(name)[:200], len(lines)))
    fam = collections.Counter()
    for op, n in c.items():
        for key in SHOW:
            if op.startswith(key):
                fam[op] += n
    print("   " + ", ".join("%s x%d" % kv for kv in sorted(fam.items())))
    seen = set()
    for a, t in lines:
        toks = t.split()
        op = toks[1] if toks[0].startswith('@') and len(toks) > 1 else toks[0]
        for key in ("UTCHMMA", "UTCBAR", "LDTM", "LDGSTS", "REDG", "RED.", "FFMA2", "CCTL", "SYNCS.PHASECHK", "UTCATOMSWS"):
            if op.startswith(key) and key not in seen:
                seen.add(key)
                print("   /*%04x*/ %s" % (a, t))
    print()
