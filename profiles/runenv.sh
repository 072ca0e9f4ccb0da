#!/bin/bash
# bench.py under different environment settings: profiles/runenv.sh VAR v1 v2 ...
var=$1; shift
for v in "$@"; do
  echo "== $var=$v"; env $var=$v python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value %.0f e2e %.0f kernel_ms %.3f e2e_ms %.3f' % (d['value'], d['e2e']['value'], d['roofline']['kernel_ms'], d['e2e']['ms_per_step']), d['kernel'], d['check'])"
done
