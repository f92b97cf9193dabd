#!/bin/bash
# developer helper (GPU box): the round's committed evidence -- plain bench first, then launch list, then full captures
set -x
python bench.py --steps 50 --warmup 5 --skip-cpu --skip-extra > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 50 --warmup 5 --skip-cpu --skip-extra > gpurun_out/prof_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_tma_kernel -s 20 -c 3 -o gpurun_out/r2_scan_tma \
    python bench.py --steps 50 --warmup 5 --skip-cpu --skip-extra > gpurun_out/prof_ncu2.log 2>&1
python tools/batch_bench.py > gpurun_out/c3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_topk -s 7 -c 1 -o gpurun_out/r2_c3_gemm \
    python tools/batch_bench.py > gpurun_out/prof_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r2_launches.csv
