#!/bin/bash
# --set full capture of the large-scan kernel on one C4 wave, with the per-instruction source page: profiles/gpu_ncu_l2.sh <tag>
tag=$1; out=gpurun_out/$tag; mkdir -p $out
CCICP_CHUNK_PAIRS=100000 timeout 900 ncu --set full --clock-control none --import-source on -k regex:icp_pair_l2_kernel -s 1 -c 1 -f -o $out/large_full \
    python profiles/bench_configs.py c4 --pairs 148 > $out/ncu_large.log 2>&1
ncu -i $out/large_full.ncu-rep --page raw --csv > $out/large_full_raw.csv 2>/dev/null
ncu -i $out/large_full.ncu-rep --page source --csv --print-source sass > $out/large_full_sass.csv 2>/dev/null
rm -f $out/large_full.ncu-rep
CCICP_CHUNK_PAIRS=100000 python profiles/bench_configs.py c4 --pairs 148 2>&1 | tail -1
python - <<PY
import csv
r=list(csv.reader(open("$out/large_full_raw.csv"))); d=dict(zip(r[0],r[2]))
for k in d:
    if k.startswith(("gpu__time_duration.sum","smsp__inst_executed.sum","smsp__issue_active.avg.pct","smsp__thread_inst_executed_per_inst","smsp__average_warps_issue_stalled","l1tex__t_sector_hit_rate","lts__t_sector_hit_rate")) : print(k, d[k])
PY
