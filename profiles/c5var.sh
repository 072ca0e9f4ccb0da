#!/bin/bash
# A/B of library variants on C5 (256 pairs x 256 seeds, best-of on device) and C2: profiles/c5var.sh v_a v_b ...
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib timeout 300 python profiles/bench_configs.py c5 2>&1 | tail -1 | cut -c1-200
done
