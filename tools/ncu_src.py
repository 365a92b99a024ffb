#!/usr/bin/env python
"""Stall samples of an .ncu-rep's SASS source page: totals by reason, by 100-instruction region, top instructions.
usage: ncu_src.py rep [top]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[1], rows[2:]
isrc, isamp, iex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = collections.Counter()
for r in data:
    for i in stall:
        tot[hdr[i]] += int(r[i] or 0)
n = sum(int(r[isamp] or 0) for r in data)
print("samples", n, "instructions", len(data), "executed", sum(int(r[iex] or 0) for r in data))
print(", ".join("%s %d" % (k[6:], v) for k, v in tot.most_common(9)))
for b in range(0, len(data), 100):
    seg = data[b:b + 100]
    s = sum(int(r[isamp] or 0) for r in seg)
    ex = sum(int(r[iex] or 0) for r in seg)
    ops = collections.Counter((r[isrc].split()[1] if r[isrc].startswith('@') else r[isrc].split()[0]) for r in seg if r[isrc].strip())
    if s > n / 150:
        print("%5d %6d (%4.1f%%) ex=%9d %s" % (b, s, 100.0 * s / n, ex, " ".join("%s:%d" % kv for kv in ops.most_common(6))))
best = sorted(((int(r[isamp] or 0), k) for k, r in enumerate(data)), reverse=True)[:top]
for s, k in best:
    r = data[k]
    t2 = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:2]
    print("%6d %4.1f%% #%5d ex=%8d %-64s %s" % (s, 100.0 * s / n, k, int(r[iex] or 0), r[isrc][:64], t2))
