"""C3 (loop-closure pairs: +-2 m / +-0.3 rad offsets, init = truth + N(0.3 m, 0.05 rad)) device-resident timing per
iteration count, beside C2 pairs of the same count.  usage: c3probe.py [pairs]"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
P = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda", 0)
ctx = capi.Context(0)
st = torch.cuda.Stream(device=dev)
for name, cfg, seed in (("C2", 1, 1000), ("C3", 3, 10**6)):
    b = scangen.generate(seed, P, config=cfg)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
    d = [t(b.model_xy, np.float64), t(b.model_off, np.int32), t(b.scene_xy, np.float64), t(b.scene_off, np.int32), t(b.init_Rt(), np.float64)]
    d_out = torch.zeros(P * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    hint = capi.max_cells(b.model_off)
    for iters, flags in ((1, 0), (2, 0), (5, 0), (10, 0), (30, 0), (30, capi.F_NO_TIE_REPLAY)):
        prm = capi.fixed_iter_params(iters, flags=flags)
        ms = []
        for _ in range(4):
            ctx.register_batch_device(P, 1, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), d[4].data_ptr(), prm,
                                      d_out.data_ptr(), 0, b.max_model, b.max_scene, st.cuda_stream, max_cells_hint=hint)
            ctx.sync(st.cuda_stream)
            ms.append(ctx.launch_info().last_kernel_ms)
        out = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=capi.RESULT_DTYPE)
        print("%s %5d pairs  iters %2d flags %2d: %8.3f ms  (%.0f pairs/s)  mean corr %.0f  replay pairs %d" % (
            name, P, iters, flags, min(ms[1:]), P / min(ms[1:]) * 1e3, out["num_corr"].mean(), int((out["status"] & capi.ST_TIE_REPLAY != 0).sum())), flush=True)
