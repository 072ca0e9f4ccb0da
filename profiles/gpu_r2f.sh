#!/bin/bash
# r2f: L1 size against tie-replay staging (CCICP_STAGE_KB) for the two large-scan variants, C4 512 pairs
out=gpurun_out/r2f; mkdir -p $out
for nc in 1 0; do for kb in 200 96 48 24; do
  echo "== NO_CLUSTER=$nc STAGE_KB=$kb"; CCICP_NO_CLUSTER=$nc CCICP_STAGE_KB=$kb timeout 300 python profiles/bench_configs.py c4 --pairs 512 | cut -c1-200
done; done > $out/stage_kb.log 2>&1
cat $out/stage_kb.log
