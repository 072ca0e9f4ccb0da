"""Timeline of one host-buffer call (C2, pinned buffers): H2D span, kernel span, D2H span as the library's events see them."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
n = 4096
b = scangen.generate(1000, n, pinned=capi.pinned_array)
prm = capi.fixed_iter_params(30)
ctx = capi.Context(0)
init = capi.pinned_array((n, 6), np.float64); init[:] = b.init_Rt()
out = capi.pinned_array((n,), capi.RESULT_DTYPE)
for rep in range(6):
    t0 = time.perf_counter()
    ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, init, prm, out=out)
    w = (time.perf_counter() - t0) * 1e3
    li = ctx.launch_info()
    print("call %.3f ms | h2d span %.3f | kernel span %.3f | d2h %.3f | launches %d" % (w, li.last_h2d_ms, li.last_kernel_ms, li.last_d2h_ms, li.n_launches), flush=True)
bytes_in = b.model_xy.nbytes + b.scene_xy.nbytes
print("input %.1f MB -> %.1f GB/s over the h2d span" % (bytes_in / 1e6, bytes_in / 1e6 / li.last_h2d_ms))
