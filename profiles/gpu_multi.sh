#!/bin/bash
# multi-GPU round: N = number of GPUs of the box (2 or 8).  C++ multi-device demo (NCCL gather), h2d ceiling, bench scaling.
N=$1; out=gpurun_out/r2m_$N; mkdir -p $out
nvidia-smi -L > $out/gpus.txt; nproc > $out/nproc.txt
devs=$(seq -s, 0 $((N-1)))
tests/host/multi_demo_b200 $devs 4096 > $out/multi_demo.log 2>&1; echo "rc=$?" >> $out/multi_demo.log; cat $out/multi_demo.log
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider > $out/pytest_multi.log 2>&1; tail -3 $out/pytest_multi.log
for n in 1 2 4 8; do
  [ $n -gt $N ] && break
  if [ $n -eq 1 ]; then python profiles/h2d_ceiling.py; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 profiles/h2d_ceiling.py; fi
done > $out/h2d_ceiling.jsonl 2> $out/h2d_ceiling.err
cat $out/h2d_ceiling.jsonl
for n in ${BENCH_NS:-2 4 8}; do
  [ $n -gt $N ] && break
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 10 --warmup 3 > $out/bench_${n}gpu.json 2> $out/bench_${n}gpu.err
  python -c "
import json,sys
d=json.loads(open('$out/bench_${n}gpu.json').read().strip().splitlines()[-1])
print('N=$n value %.0f e2e %.0f ms %.3f kernel %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms']), 'c3:', {k: d['configs'].get('c3', {}).get(k) for k in ('value','ms_per_step','kernel_ms')}, d['configs'].get('c3', {}).get('e2e', {}).get('value'))
" 2>&1 | tail -2
done
python profiles/multi_bench.py $devs > $out/multi_abi_bench.json 2> $out/multi_abi_bench.err; cat $out/multi_abi_bench.json
