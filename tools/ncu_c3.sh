#!/bin/bash
# developer helper (GPU box): launch list of one C3 batch + timing without the profiler
set -e
python tools/batch_bench.py > gpurun_out/c3_plain.log 2>&1 && cat gpurun_out/c3_plain.log &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 15 -c 12 --csv --log-file gpurun_out/c3_launches.csv python tools/batch_bench.py > gpurun_out/c3_ncu.log 2>&1
python - <<'PY'
import csv
rows = list(csv.reader(l for l in open("gpurun_out/c3_launches.csv") if l.startswith('"')))
h = rows[0]
for r in rows[1:]:
    if r[h.index("Metric Name")] == "gpu__time_duration.sum":
        print(r[h.index("Kernel Name")][:70], r[h.index("Metric Value")], r[h.index("Metric Unit")])
PY
