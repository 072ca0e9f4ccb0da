#!/bin/bash
# A/B of library variants: profiles/runvar.sh v_old v_g128 ...   (each slam_b200/lib/<name>.so; "default" = the product build)
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['kernel_ms'], d['kernel'], d['check'])"
done
