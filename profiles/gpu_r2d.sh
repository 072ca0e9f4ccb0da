#!/bin/bash
# r2d: cluster kernel parity + timing against the L2 variant, kernel variant A/B
out=gpurun_out/r2d; mkdir -p $out
timeout 900 python -m pytest tests/test_large_gpu.py tests/test_configs_gpu.py tests/test_pipeline_gpu.py tests/test_fuzz_gpu.py -m gpu -q -x -p no:cacheprovider > $out/pytest_large.log 2>&1; echo "rc=$?" >> $out/pytest_large.log
tail -15 $out/pytest_large.log
for v in 0 1; do echo "== CCICP_NO_CLUSTER=$v"; CCICP_NO_CLUSTER=$v timeout 600 python profiles/bench_configs.py c4 --pairs 1024; done > $out/c4_cluster_vs_l2.log 2>&1
cat $out/c4_cluster_vs_l2.log
bash profiles/runvar.sh default v_solveall v_mb3 v_row2 v_flat > $out/variants.log 2>&1
cat $out/variants.log
