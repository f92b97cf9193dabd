#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (argv[1]); argv[2] = the command it came from."""
import collections, csv, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]; K, M, V, U = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    if r[M] != "gpu__time_duration.sum": continue
    us = float(r[V].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[U], 1e-3)
    tot[r[K]] += us; cnt[r[K]] += 1
allus = sum(tot.values())
cmd = sys.argv[2] if len(sys.argv) > 2 else "?"
print(f"launch list of `{cmd}` under ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
for k, us in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{us:12.1f} us {100 * us / allus:6.2f} % {cnt[k]:6d} launches {us / cnt[k]:10.2f} us avg  {k[:70]}")
