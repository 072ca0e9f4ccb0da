#!/bin/bash
# r2e: why is the cluster kernel slower than the L2 kernel on C4?  cluster sizes 2 / 4 / 8, one ncu --set full capture
out=gpurun_out/r2e; mkdir -p $out
for v in v_cl2 default v_cl8; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib timeout 600 python profiles/bench_configs.py c4 --pairs 512
done > $out/cluster_sizes.log 2>&1
cat $out/cluster_sizes.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:icp_pair_cluster -s 1 -c 1 -f -o $out/cluster_full \
    python profiles/bench_configs.py c4 --pairs 132 > $out/ncu_cluster.log 2>&1
ncu -i $out/cluster_full.ncu-rep --page raw --csv > $out/cluster_full_raw.csv 2>/dev/null
ncu -i $out/cluster_full.ncu-rep --page source --csv --print-source sass > $out/cluster_full_sass.csv 2>/dev/null
sha256sum slam_b200/lib/libccicp2d_b200.so | cut -d' ' -f1 > $out/lib_sha256.txt
tail -3 $out/ncu_cluster.log
