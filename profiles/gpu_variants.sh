#!/bin/bash
# A/B of kernel variants built by profiles/buildvar.sh: bench line of each, then the parity tests that exercise the search
# through each variant library.   usage: profiles/gpu_variants.sh <tag> v_a v_b ...
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
bash profiles/runvar.sh default "$@" default > $out/variants.log 2>&1
for v in "$@"; do
  echo "== parity through $v" >> $out/variants.log
  CCICP_LIB=$PWD/slam_b200/lib/$v.so timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_fuzz_gpu.py tests/test_configs_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3 >> $out/variants.log
done
cat $out/variants.log
