#!/usr/bin/env python
"""Instructions executed and stall samples per CUDA source line of the first kernel in an .ncu-rep (-lineinfo build)."""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 50
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
fname, out, hdr = "", [], None
for r in csv.reader(raw.splitlines()):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and r[0].isdigit():
        N, X = hdr.index("# Samples"), hdr.index("Instructions Executed")
        try: out.append((float(r[X]), float(r[N]), fname, int(r[0]), r[1].strip()))
        except ValueError: pass
ti, ts = sum(o[0] for o in out), sum(o[1] for o in out)
print(f"total warp instructions {ti:.0f}, samples {ts:.0f}")
for x, n, f, ln, src in sorted(out, reverse=True)[:topn]:
    print(f"{100*x/ti:5.1f}% inst {100*n/ts:5.1f}% smp  {f}:{ln:<4d} {src[:110]}")
