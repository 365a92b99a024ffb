#!/usr/bin/env python
"""Print the SASS of one kernel of a .so (substring match on the mangled name) and an opcode
histogram of an address range.  usage: sass_fn.py lib.so substr [lo hi]   (hex addresses)"""
import collections
import re
import subprocess
import sys

so, sub = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
blocks = txt.split("Function : ")
blk = [b for b in blocks if b.split("\n", 1)[0].find(sub) >= 0][0]
lines = []
for ln in blk.split("\n"):
    m = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', ln)
    if m:
        lines.append((int(m.group(1), 16), re.sub(r'\s+', ' ', m.group(2)).strip()))
lo = int(sys.argv[3], 16) if len(sys.argv) > 3 else 0
hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else 1 << 30
sel = [(a, t) for a, t in lines if lo <= a <= hi]
c = collections.Counter()
for a, t in sel:
    toks = t.split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    c[op.split('.')[0]] += 1
print(blk.split("\n", 1)[0], "instructions:", len(sel))
print(c.most_common(30))
if len(sys.argv) > 3:
    for a, t in sel:
        print(hex(a), t)
