#!/bin/bash
# ncu captures of the kernels that are not on the bench.py line: L2 variant (C4), classification + radix sort, LUM.
out=gpurun_out/${1:-extra}; mkdir -p $out
python profiles/run_c4.py > $out/c4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:icp_pair_l2 -s 2 -c 1 -f -o $out/l2_full python profiles/run_c4.py > $out/ncu_l2.log 2>&1
ncu -i $out/l2_full.ncu-rep --page raw --csv > $out/l2_raw.csv 2>/dev/null
python profiles/bench_configs.py classify lum > $out/extra_bench.log 2>&1 && \
ncu --set full --clock-control none -k regex:"classify|rs_|lum_" -s 22 -c 30 -f -o $out/cls_full python profiles/bench_configs.py classify lum > $out/ncu_cls.log 2>&1
ncu -i $out/cls_full.ncu-rep --page raw --csv > $out/cls_raw.csv 2>/dev/null
cat $out/c4.log; cat $out/extra_bench.log | cut -c1-300
