#!/bin/bash
# ncu --set full capture of one classifyPoints call (17 launches) after one warm-up call.
out=gpurun_out/${1:-cls}; mkdir -p $out
python profiles/bench_configs.py classify > $out/classify_bench.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"classify|seg_|scan_|plan_" -s 17 -c 17 -f -o $out/cls_full python profiles/bench_configs.py classify > $out/ncu_cls.log 2>&1
ncu -i $out/cls_full.ncu-rep --page raw --csv > $out/cls_raw.csv 2>/dev/null
cat $out/classify_bench.log | cut -c1-400
