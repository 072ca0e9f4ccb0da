#!/bin/bash
# e2e leg under several host-path chunk settings: profiles/runenv2.sh "CHUNK_PAIRS RAMP" ...
for cfg in "$@"; do
  set -- $cfg
  echo "== CCICP_CHUNK_PAIRS=$1 CCICP_CHUNK_RAMP=$2"
  CCICP_CHUNK_PAIRS=$1 CCICP_CHUNK_RAMP=$2 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value %.0f e2e %.0f kernel_ms %.3f e2e_ms %.3f' % (d['value'], d['e2e']['value'], d['roofline']['kernel_ms'], d['e2e']['ms_per_step']))"
done
