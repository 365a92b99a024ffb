#!/usr/bin/env python
"""Condense `nvcc -Xptxas -v` output: one line per kernel (registers, spills, smem).
usage: python -m kinet_b200.build --force -v 2>&1 | python tools/regs.py [filter]"""
import re
import subprocess
import sys

flt = sys.argv[1] if len(sys.argv) > 1 else ""
name = None
spill = ""
for ln in sys.stdin:
    m = re.search(r"Compiling entry function '(\S+)'", ln)
    if m:
        try:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        except Exception:
            name = m.group(1)
        name = re.sub(r"\(.*", "", name).replace("void msda::", "")
        continue
    if "spill" in ln:
        spill = ln.strip().replace("bytes", "B")
    m = re.search(r"Used (\d+) registers(?:, used \d+ barriers)?(.*)", ln)
    if m and name:
        if flt in name:
            sp = "" if " 0 B spill stores, 0 B spill loads" in spill.replace("bytes", "B") else "  SPILL: " + spill
            print("%-70s regs %3s %s%s" % (name[:70], m.group(1), m.group(2).strip()[:60], sp))
        name = None
    if "error" in ln or "warning" in ln:
        print(ln.rstrip())
