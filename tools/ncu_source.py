#!/usr/bin/env python
"""Top stalled SASS instructions from an .ncu-rep (needs --import-source on / -lineinfo)."""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
# first kernel only
start = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
end = next((i for i in range(start + 1, len(rows)) if rows[i] and rows[i][0] == 'Kernel Name'), len(rows))
hdr = rows[start]
S, N, X = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for k, r in enumerate(rows[start + 1:end]):
    try: data.append((float(r[N]), k, r))
    except (ValueError, IndexError): pass
tot = sum(d[0] for d in data)
print('total samples', tot, 'instructions', len(data))
for s, k, r in sorted(data, key=lambda x: -x[0])[:topn]:
    st = sorted(((float(r[i] or 0), h[6:]) for i, h in stall_cols), reverse=True)[:2]
    print(f"{s:7.0f} {100*s/tot:5.1f}% #{k:5d} exec={r[X]:>9s} {r[S].strip()[:70]:70s} {st[0][1]}:{st[0][0]:.0f} {st[1][1]}:{st[1][0]:.0f}")
