"""North_star extra (no reference counterpart, parity unpinned by construction): per-point curvature / normal features
(ccicp_scan_features) and curvature-gated matching (CCICP_F_CURV_GATE) against their CPU statement oracle/curv_port.py.
With the flag off nothing changes -- every parity test of the reference path runs without it."""
import numpy as np
import pytest

from oracle import curv_port as cv
from slam_b200 import capi, scangen
from helpers import csr_from_pairs, gpu_lockstep_corr

pytestmark = pytest.mark.gpu


def test_features_match_the_cpu_statement(ctx):
    b = scangen.generate(300, 3)
    for xy, off in ((b.model_xy, b.model_off), (b.scene_xy, b.scene_off)):
        for k, radius in ((5, 0.5), (2, 0.08), (8, 3.0)):
            nrm, cur = ctx.scan_features(xy, off, k, radius)
            wn, wc = cv.features(xy, off, k, radius)
            assert np.array_equal(np.isnan(cur), np.isnan(wc))
            np.testing.assert_allclose(cur, wc, rtol=0, atol=2e-7, equal_nan=True)
            np.testing.assert_allclose(nrm, wn, rtol=0, atol=2e-6)
    nrm, cur = ctx.scan_features(b.model_xy, b.model_off)
    ok = ~np.isnan(cur)
    assert ok.mean() > 0.9 and np.median(cur[ok]) < 0.05 and cur[ok].max() > 0.2         # walls are flat (1 cm range noise), corners are not
    assert np.abs(np.linalg.norm(nrm[ok & (np.abs(nrm).sum(1) > 0)], axis=1) - 1).max() < 1e-6


@pytest.mark.parametrize("tol", [0.002, 0.02, 1.0])
def test_curvature_gated_fit_matches_the_cpu_statement(ctx, tol):
    b = scangen.generate(310, 3)
    prm = capi.fixed_iter_params(8, flags=capi.F_CURV_GATE)
    prm.curv_tol = tol
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    plain = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), capi.fixed_iter_params(8))
    so = b.scene_off
    for p in range(3):
        want = cv.fit_curv(*b.pair(p), b.init_Rt()[p], max_iter=8, tol=tol)
        got = corr[0, so[2 * p]:so[2 * p + 2]]
        assert (got == want["corr"][-1]).all(), "pair %d: %d of %d indices differ" % (p, (got != want["corr"][-1]).sum(), len(got))
        assert res[p]["num_corr"] == want["ncorr"][-1]
        np.testing.assert_allclose(np.r_[res[p]["R"], res[p]["t"]], want["Rt"], atol=1e-9)
        if tol >= 1.0:       # curvature lies in [0, 0.5]: a tolerance of 1 gates nothing -> the plain fit's correspondences
            assert res[p]["num_corr"] == plain[p]["num_corr"]
        else:
            assert res[p]["num_corr"] <= plain[p]["num_corr"]
    assert not (res["status"] & capi.ST_TIE_REPLAY).any()


def test_curvature_gate_lockstep_and_flag_checks(ctx):
    b = scangen.generate(320, 1)
    want = cv.fit_curv(*b.pair(0), b.init_Rt()[0], max_iter=5, tol=0.01)
    corr, res = gpu_lockstep_corr(ctx, b.pair(0), want["Rt_trace"][:-1], flags=capi.F_CURV_GATE)       # default curv_tol is 0.05
    want5 = cv.fit_curv(*b.pair(0), b.init_Rt()[0], max_iter=1, tol=0.05)
    assert (corr[0] == want5["corr"][0]).all()
    prm = capi.fixed_iter_params(3, flags=capi.F_CURV_GATE | capi.F_EXACT_SUMS)
    with pytest.raises(capi.CcicpError):
        ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
