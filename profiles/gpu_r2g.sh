#!/bin/bash
out=gpurun_out/r2g; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/pytest_gpu.log
tail -12 $out/pytest_gpu.log
timeout 300 python profiles/run_trim.py > $out/trim.log 2>&1; cat $out/trim.log
