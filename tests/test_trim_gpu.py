"""CCICP_F_TRIM (north_star extra: trimmed outlier rejection, no counterpart in the reference) against its CPU statement
oracle/trim_port.py: correspondence indices of the last iteration bit-exact, pose within 1e-5 m / 1e-6 rad."""
import numpy as np
import pytest

from slam_b200 import capi, scangen
from oracle import trim_port
from helpers import assert_pose_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("frac,iters", [(0.9, 30), (0.6, 10), (1.0, 5)])
def test_trimmed_registration_matches_cpu_statement(ctx, frac, iters):
    b = scangen.generate(900, 6, config=3)
    prm = capi.fixed_iter_params(iters, flags=capi.F_TRIM, trim_frac=frac)
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    so = b.scene_off
    for p in range(b.n_pairs):
        want = trim_port.fit_trimmed(*b.pair(p), b.init_Rt()[p], max_iter=iters, trim_frac=frac)
        got = corr[0, so[2 * p]:so[2 * p + 2]]
        assert (got == want["corr"]).all(), "pair %d: %d indices differ" % (p, (got != want["corr"]).sum())
        assert res[p]["num_corr"] == want["num_corr"] and res[p]["iters"] == iters
        assert_pose_close(np.r_[res[p]["R"], res[p]["t"]], want["Rt"], "pair %d" % p)


def test_trim_off_equals_default_kernel(ctx):
    b = scangen.generate(901, 8)
    a = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), capi.fixed_iter_params(30))
    t = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(),
                           capi.fixed_iter_params(30, flags=capi.F_TRIM, trim_frac=1.0))
    for f in ("R", "t", "num_corr"):
        assert np.array_equal(a[f], t[f])


def test_trim_through_the_l2_variant(ctx):
    import os
    os.environ["CCICP_FORCE_L2"] = "1"
    try:
        c = capi.Context(0)
    finally:
        del os.environ["CCICP_FORCE_L2"]
    b = scangen.generate(902, 3, config=3)
    prm = capi.fixed_iter_params(8, flags=capi.F_TRIM, trim_frac=0.75)
    res, corr = c.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    r2, c2 = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    assert c.launch_info().kernel_variant == 1
    assert (corr == c2).all() and np.array_equal(res["num_corr"], r2["num_corr"])
    c.close()


def test_invalid_trim_arguments(ctx):
    b = scangen.generate(1, 1)
    args = (b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt())
    for prm in (capi.fixed_iter_params(3, flags=capi.F_TRIM, trim_frac=0.0),
                capi.fixed_iter_params(3, flags=capi.F_TRIM, trim_frac=1.5),
                capi.fixed_iter_params(3, flags=capi.F_TRIM | capi.F_EXACT_SUMS, trim_frac=0.5)):
        with pytest.raises(capi.CcicpError):
            ctx.register_batch(*args, prm)
