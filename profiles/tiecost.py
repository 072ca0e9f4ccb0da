"""What do the tie replays of the C2 batch cost?  Same library, same batch, one device-resident launch, with and
without CCICP_F_NO_TIE_REPLAY (the flag leaves the tied queries with the grid's answer: wrong for parity, but it
isolates the replay WORK from the cost of carrying the replay CODE in the kernel)."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
n = 4096
b = scangen.generate(1000, n, config=1)
dev = torch.device("cuda", 0)
t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
d_mxy, d_sxy = t(b.model_xy, np.float64), t(b.scene_xy, np.float64)
d_moff, d_soff = t(b.model_off, np.int32), t(b.scene_off, np.int32)
d_init = t(b.init_Rt(), np.float64)
d_out = torch.zeros(n * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
ctx = capi.Context(0)
hint = capi.max_cells(b.model_off)
for name, flags in (("replay", 0), ("no replay", capi.F_NO_TIE_REPLAY)) * 2:
    prm = capi.fixed_iter_params(30, flags=flags)
    ms = []
    for _ in range(8):
        ctx.register_batch_device(n, 1, d_mxy.data_ptr(), d_moff.data_ptr(), d_sxy.data_ptr(), d_soff.data_ptr(),
                                  d_init.data_ptr(), prm, d_out.data_ptr(), 0, b.max_model, b.max_scene, 0, max_cells_hint=hint)
        torch.cuda.synchronize()
        ms.append(ctx.launch_info().last_kernel_ms)
    out = d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    print("%-10s kernel ms %.3f (min %.3f)  tied pairs %d replays %d" % (name, np.mean(ms[3:]), np.min(ms[3:]), int((out["n_tie_replays"] > 0).sum()), int(out["n_tie_replays"].sum())), flush=True)
