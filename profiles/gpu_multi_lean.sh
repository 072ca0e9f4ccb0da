#!/bin/bash
# lean multi-GPU check of the final library: C++ multi-device demo, the multi-GPU tests, bench.py at N GPUs, the multi-device C ABI bench
N=$1; tag=${2:-r2n}; out=gpurun_out/${tag}_$N; mkdir -p $out
devs=$(seq -s, 0 $((N-1)))
tests/host/multi_demo_b200 $devs 4096 > $out/multi_demo.log 2>&1; echo "rc=$?" >> $out/multi_demo.log; cat $out/multi_demo.log
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider > $out/pytest_multi.log 2>&1; tail -2 $out/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > $out/bench_${N}gpu.json 2> $out/bench_${N}gpu.err
python -c "
import json
d=json.loads(open('$out/bench_${N}gpu.json').read().strip().splitlines()[-1])
print('N=$N value %.0f e2e %.0f ms %.3f kernel %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms']), 'c3:', {k: d['configs'].get('c3', {}).get(k) for k in ('value','ms_per_step','kernel_ms')}, d['configs'].get('c3', {}).get('e2e', {}).get('value'))
" 2>&1 | tail -2
python profiles/multi_bench.py $devs > $out/multi_abi_bench.json 2> $out/multi_abi_bench.err; tail -1 $out/multi_abi_bench.json | cut -c1-600
