#!/bin/bash
# A/B of library variants on the C4 wave (148 pairs x 32768 points, L2 variant): profiles/c4var.sh v_a v_b ...
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib timeout 200 python profiles/c4probe.py 200 2>&1 | sed -n 2,3p
done
