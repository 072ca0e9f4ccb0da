"""One C4 wave (148 pairs x 32768 points, L2 variant) -- profiling target."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
ctx = capi.Context(0)
b = scangen.generate(5000, 148, config=4)
prm = capi.fixed_iter_params(30)
for _ in range(3):
    ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    print(ctx.launch_info().last_kernel_ms)
