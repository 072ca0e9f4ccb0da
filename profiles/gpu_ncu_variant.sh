#!/bin/bash
# one --set full capture of icp_pair_kernel through a variant library: profiles/gpu_ncu_variant.sh <tag> <variant>
tag=$1; v=$2; out=gpurun_out/$tag; mkdir -p $out
export CCICP_LIB=$PWD/slam_b200/lib/$v.so
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs > $out/bench.json 2> $out/bench.err || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:icp_pair_kernel -s 3 -c 1 -f -o $out/full \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs > $out/ncu.log 2>&1
ncu -i $out/full.ncu-rep --page raw --csv > $out/full_raw.csv 2>/dev/null
ncu -i $out/full.ncu-rep --page source --csv --print-source sass > $out/full_sass.csv 2>/dev/null
rm -f $out/full.ncu-rep
python - <<PY
import csv
r=list(csv.reader(open("$out/full_raw.csv"))); d=dict(zip(r[0],r[2]))
for k in d:
    if k.startswith(("gpu__time_duration.sum","smsp__inst_executed.sum","smsp__issue_active.avg.pct","smsp__thread_inst_executed_per_inst","smsp__average_warp_latency_issue_stalled","smsp__average_warps_issue_stalled")) : print(k, d[k])
PY
