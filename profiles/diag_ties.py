import sys, time, numpy as np
sys.path.insert(0, '.')
from slam_b200 import capi, scangen
ctx = capi.Context(0)
b = scangen.generate(1000, 4096)
for flags in (0, capi.F_NO_TIE_REPLAY):
    prm = capi.fixed_iter_params(30, flags=flags)
    for _ in range(3):
        res = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
        print('flags', flags, 'kernel_ms', ctx.launch_info().last_kernel_ms)
    tie = (res['status'] & capi.ST_TIE_REPLAY) != 0
    print('pairs with tie replay', tie.sum(), 'replays', res['n_tie_replays'][tie])
