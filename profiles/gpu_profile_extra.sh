#!/bin/bash
# ncu captures of the kernels that are not on the bench.py line: L2 variant (C4), LUM edge kernel.
# (classifyPoints has its own script: gpu_profile_classify.sh)
out=gpurun_out/${1:-extra}; mkdir -p $out
python profiles/run_c4.py > $out/c4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:icp_pair_l2 -s 2 -c 1 -f -o $out/l2_full python profiles/run_c4.py > $out/ncu_l2.log 2>&1
ncu -i $out/l2_full.ncu-rep --page raw --csv > $out/l2_raw.csv 2>/dev/null
python profiles/bench_configs.py lum > $out/lum_bench.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lum_edge -s 2 -c 1 -f -o $out/lum_full python profiles/bench_configs.py lum > $out/ncu_lum.log 2>&1
ncu -i $out/lum_full.ncu-rep --page raw --csv > $out/lum_raw.csv 2>/dev/null
cat $out/c4.log; cat $out/lum_bench.log | cut -c1-300
