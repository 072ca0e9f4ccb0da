"""C2 kernel time against the iteration count: where along the 30 iterations the time goes."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
import torch
n = 4096
b = scangen.generate(1000, n)
dev = torch.device("cuda", 0)
t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
d = [t(b.model_xy, np.float64), t(b.model_off, np.int32), t(b.scene_xy, np.float64), t(b.scene_off, np.int32), t(b.init_Rt(), np.float64)]
d_out = torch.zeros(n * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
ctx = capi.Context(0)
mm = int(np.diff(b.model_off[::2]).max()); ms = int(np.diff(b.scene_off[::2]).max())
hint = capi.max_cells(b.model_off)
truth = t(scangen.pose_to_Rt(b.truth), np.float64)
for it, ini, name in [(0, d[4], "init"), (1, d[4], "init"), (1, truth, "truth"), (2, d[4], "init"), (2, truth, "truth"), (30, d[4], "init"), (30, truth, "truth")]:
    prm = capi.fixed_iter_params(it, flags=capi.F_NO_EXACT_RETRY)
    ks = []
    for rep in range(4):
        ctx.register_batch_device(n, 1, *[x.data_ptr() for x in d[:4]], ini.data_ptr(), prm, d_out.data_ptr(), 0, mm, ms, None, max_cells_hint=hint)
        ctx.sync(); torch.cuda.synchronize()
        ks.append(ctx.launch_info().last_kernel_ms)
    k = min(ks[1:])
    print("iters %2d  start=%s kernel %.3f ms" % (it, name, k), flush=True)
