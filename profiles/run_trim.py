"""Throughput of the CCICP_F_TRIM extra on the C2 batch (device time of the host-buffer call with pinned arrays)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_configs import run
ctx = capi.Context(0)
b = scangen.generate(1000, 4096, config=1, pinned=capi.pinned_array)
run(ctx, "C2 default", b, capi.fixed_iter_params(30))
for f in (1.0, 0.9, 0.7):
    run(ctx, "C2 CCICP_F_TRIM trim_frac %.1f" % f, b, capi.fixed_iter_params(30, flags=capi.F_TRIM, trim_frac=f))
