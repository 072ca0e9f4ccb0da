"""C4 (32768-point pairs, L2 variant) time breakdown: iterations, tie replay, staging size, first-iteration search."""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
b = scangen.generate(5000, 148, config=4)
truth = scangen.pose_to_Rt(b.truth)
NR = capi.F_NO_TIE_REPLAY
def run(ctx, name, prm, init=None):
    init = b.init_Rt() if init is None else init
    for _ in range(2):
        out = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, init, prm)
    print("%-28s %8.2f ms  replays %d in %d pairs" % (name, ctx.launch_info().last_kernel_ms, int(out['n_tie_replays'].sum()), int((out['n_tie_replays'] > 0).sum())), flush=True)
for kb in [int(x) for x in (sys.argv[1:] or ["96"])]:
    os.environ["CCICP_STAGE_KB"] = str(kb)
    ctx = capi.Context(0)
    print("== stage %d KB" % kb)
    run(ctx, '30it', capi.fixed_iter_params(30))
    run(ctx, '30it noreplay', capi.fixed_iter_params(30, flags=NR))
    run(ctx, '1it noreplay', capi.fixed_iter_params(1, flags=NR))
    run(ctx, '2it noreplay', capi.fixed_iter_params(2, flags=NR))
    run(ctx, '1it noreplay init=truth', capi.fixed_iter_params(1, flags=NR), truth)
    run(ctx, '30it noreplay init=truth', capi.fixed_iter_params(30, flags=NR), truth)
    ctx.close()
