#!/bin/bash
# A/B of variant libraries on C2 + C5 (bench.py) and one C4 wave, then the parity tests through the first variant
# usage: profiles/gpu_ab.sh <tag> v_a v_b ...
tag=$1; shift; out=gpurun_out/$tag; mkdir -p $out
bash profiles/runvar_cfg.sh "$@" > $out/ab.log 2>&1
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== c4 $v" >> $out/ab.log; CCICP_LIB=$lib python profiles/bench_configs.py c4 --pairs 148 2>&1 | tail -1 | cut -c1-120 >> $out/ab.log
done
v=$1; lib=$PWD/slam_b200/lib/$v.so
echo "== parity through $v" >> $out/ab.log
CCICP_LIB=$lib timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_fuzz_gpu.py tests/test_configs_gpu.py tests/test_large_gpu.py tests/test_curv_gpu.py tests/test_trim_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -2 >> $out/ab.log
cat $out/ab.log
