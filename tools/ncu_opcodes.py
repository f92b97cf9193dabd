#!/usr/bin/env python
"""Executed warp instructions per opcode for the first kernel of an .ncu-rep; argv[2] = divisor (e.g. row pairs scanned)."""
import csv, subprocess, collections, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = list(csv.reader(raw.splitlines()))
start = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[start]; S = hdr.index('Source'); X = hdr.index('Instructions Executed'); N = hdr.index('# Samples')
ops = collections.Counter(); smp = collections.Counter(); tot = 0
for r in rows[start + 1:]:
    try: x = float(r[X])
    except (ValueError, IndexError): continue
    t = r[S].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    ops[op] += x; smp[op] += float(r[N]); tot += x
print(f"total {tot:.0f} = {tot / div:.1f} per unit")
for op, x in ops.most_common(int(sys.argv[3]) if len(sys.argv) > 3 else 30):
    print(f"{op:12s} {x / div:9.1f}  {100 * x / tot:5.1f}%  samples {smp[op]:.0f}")
