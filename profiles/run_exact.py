"""One C2 batch in CCICP_F_EXACT_SUMS mode (profiling target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
ctx = capi.Context(0)
b = scangen.generate(1000, 1024)
prm = capi.fixed_iter_params(30, flags=capi.F_EXACT_SUMS)
for _ in range(3):
    ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    print(ctx.launch_info().last_kernel_ms)
