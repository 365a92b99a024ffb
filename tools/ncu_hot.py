#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: instructions executed by opcode and the hottest
SASS lines.  usage: ncu_hot.py src.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hdr = rows[1]
ci, cs, csamp = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
tot = 0
byop = collections.Counter()
samp = collections.Counter()
for r in rows[2:]:
    try:
        n = int(r[ci])
    except Exception:
        continue
    src = r[cs].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '?')
    op = op.split('.')[0]
    tot += n
    byop[op] += n
    samp[op] += int(r[csamp] or 0)
print("total warp instructions", tot)
ts = sum(samp.values()) or 1
for op, n in byop.most_common(28):
    print("%-10s %12d %5.1f%%   samples %5.1f%%" % (op, n, 100.0 * n / tot, 100.0 * samp[op] / ts))
if top:
    print("--- hottest lines by samples")
    body = [r for r in rows[2:] if r[csamp].isdigit()]
    for r in sorted(body, key=lambda r: -int(r[csamp]))[:top]:
        print(r[csamp].rjust(7), r[ci].rjust(10), r[cs])
