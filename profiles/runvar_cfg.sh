#!/bin/bash
# like runvar.sh but with the C5 sub-record (C4 cut to one wave): profiles/runvar_cfg.sh v_a v_b ...
for v in "$@"; do
  lib=$PWD/slam_b200/lib/$v.so; [ "$v" = default ] && lib=$PWD/slam_b200/lib/libccicp2d_b200.so
  echo "== $v"; CCICP_LIB=$lib python bench.py --steps 5 --warmup 3 --no-cpu-baseline --c4-pairs 148 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c2 %.0f e2e %.0f kernel_ms %.4f  c5 %.0f (%.1f ms)' % (d['value'], d['e2e']['value'], d['roofline']['kernel_ms'], d['configs']['c5']['value'], d['configs']['c5']['kernel_ms']), d['check'])"
done
