#!/bin/bash
# --set full capture of the multi-seed kernel on the full C5 batch (1024 pairs x 256 seeds): profiles/gpu_ncu_seeds.sh <tag>
tag=$1; out=gpurun_out/$tag; mkdir -p $out
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:icp_pair_seeds_kernel -s 1 -c 1 -f -o $out/seeds_full \
    python profiles/bench_configs.py c5 --pairs 1024 > $out/ncu_seeds.log 2>&1
ncu -i $out/seeds_full.ncu-rep --page raw --csv > $out/seeds_full_raw.csv 2>/dev/null
rm -f $out/seeds_full.ncu-rep
sha256sum slam_b200/lib/libccicp2d_b200.so | cut -d" " -f1 > $out/lib_sha256.txt
tail -3 $out/ncu_seeds.log | cut -c1-300
