#!/bin/bash
# C4 wave through variant libraries + the large-scan parity tests: profiles/runvar_c4.sh v_a v_b ...
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib python profiles/bench_configs.py c4 --pairs 148 2>&1 | tail -1 | cut -c1-200
  [ "$v" != default ] && CCICP_LIB=$lib timeout 900 python -m pytest tests/test_large_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -1
done
