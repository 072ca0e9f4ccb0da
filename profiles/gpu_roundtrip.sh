#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (both arms), then -- only after the un-profiled bench exited 0 --
# the ncu launch list and one --set full capture of icp_pair_kernel (and, with "l2", of the large-scan kernel).
# Outputs land in gpurun_out/<tag>/.   usage: profiles/gpu_roundtrip.sh <tag> [skip-tests] [l2]
tag=${1:-r2}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
nproc > $out/nproc.txt
profiles/fp32_peak > $out/fp32_unfused_peak.json 2> $out/fp32.err
if [ "$2" != "skip-tests" ]; then
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > $out/pytest_no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> $out/pytest_no_alloc_alone.log
  timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/pytest_gpu.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke rc=$?" >> $out/smoke.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 > $out/bench_1gpu.json 2> $out/bench_1gpu.err; rc=$?
echo "bench rc=$rc" >> $out/bench_1gpu.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_ref.json 2> $out/bench_ref.err
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs > $out/ncu_launches.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:icp_pair_kernel -s 3 -c 1 -f -o $out/icp_pair_full \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs > $out/ncu_full.log 2>&1
  ncu -i $out/icp_pair_full.ncu-rep --page raw --csv > $out/icp_pair_full_raw.csv 2>/dev/null
  ncu -i $out/icp_pair_full.ncu-rep --page source --csv --print-source sass > $out/icp_pair_full_sass.csv 2>/dev/null
  sha256sum slam_b200/lib/libccicp2d_b200.so | cut -d' ' -f1 > $out/lib_sha256.txt
  if [ "$3" == "l2" ]; then
    # (CCICP_CHUNK_PAIRS: one launch of all 148 pairs -- the host-buffer call would otherwise split 1 MB pairs into chunks of 16)
    CCICP_CHUNK_PAIRS=100000 timeout 900 ncu --set full --clock-control none --import-source on -k regex:icp_pair_l2_kernel\|icp_pair_cluster -s 1 -c 1 -f -o $out/large_full \
        python profiles/bench_configs.py c4 --pairs 148 > $out/ncu_large.log 2>&1
    ncu -i $out/large_full.ncu-rep --page raw --csv > $out/large_full_raw.csv 2>/dev/null
  fi
fi
tail -3 $out/pytest_no_alloc_alone.log 2>/dev/null; tail -3 $out/pytest_gpu.log 2>/dev/null; tail -2 $out/smoke.log 2>/dev/null; cat $out/bench_1gpu.json; tail -2 $out/bench_1gpu.err; cat $out/bench_ref.json
