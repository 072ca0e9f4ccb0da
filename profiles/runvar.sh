for v in g64 g96 g128; do
  echo "== $v"; CCICP_LIB=$PWD/slam_b200/lib/v_$v.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['kernel_ms'], d['kernel'], d['check'])"
done
