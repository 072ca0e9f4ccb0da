"""Distance evaluations per query in the first (cold) iteration of C2 pairs against later iterations (diagnostics kernel)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slam_b200 import capi, scangen
b = scangen.generate(1000, 1024)
ctx = capi.Context(0)
prev = (0, 0)
for it in (1, 2, 3, 5, 10, 30):
    ctx.diag_counts()
    ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), capi.fixed_iter_params(it, flags=capi.F_DIAG_COUNTS))
    ev, q, inl = ctx.diag_counts()
    print("iters %2d: evals/query %.2f overall; iterations %d..%d: %.2f" % (it, ev / q, prev[2] + 1 if len(prev) > 2 else 1, it, (ev - prev[0]) / max(q - prev[1], 1)))
    prev = (ev, q, it)
