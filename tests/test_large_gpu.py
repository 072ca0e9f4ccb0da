"""GPU parity of the large-scan variants of the pair kernel (working set in a global workspace): the cluster-per-pair
kernel (kernel_variant 4, opt-in with CCICP_CLUSTER=1) and the one-CTA-per-pair L2 kernel (kernel_variant 1, the default for
scans beyond one SM's shared memory, and the only one for the exact-order / trim instantiations).  Config C4 (32768-point dense
slices) and, with CCICP_FORCE_L2=1, the ordinary 1081-beam pairs routed through the same variants."""
import os

import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import assert_pose_close, gpu_lockstep_corr

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["cluster", "l2"])
def l2ctx(request):
    """Every batch through a large-scan variant: the one-CTA L2 kernel, or (CCICP_CLUSTER=1) the cluster kernel."""
    os.environ["CCICP_FORCE_L2"] = "1"
    if request.param == "cluster":
        os.environ["CCICP_CLUSTER"] = "1"
    try:
        c = capi.Context(0)
    finally:
        del os.environ["CCICP_FORCE_L2"]
        os.environ.pop("CCICP_CLUSTER", None)
    c.want_variant = 4 if request.param == "cluster" else 1
    yield c
    c.close()


@pytest.fixture()
def clctx():
    os.environ["CCICP_CLUSTER"] = "1"
    try:
        c = capi.Context(0)
    finally:
        del os.environ["CCICP_CLUSTER"]
    yield c
    c.close()


def test_forced_l2_matches_oracle_and_onchip(ctx, l2ctx, oracle):
    b = scangen.generate(1000, 24)
    prm = capi.fixed_iter_params(30, flags=capi.F_INFO)
    args = (b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    r_on, c_on = ctx.register_batch(*args, want_corr=True)
    r_l2, c_l2 = l2ctx.register_batch(*args, want_corr=True)
    assert ctx.launch_info().kernel_variant == 0 and l2ctx.launch_info().kernel_variant == l2ctx.want_variant
    assert (c_on == c_l2).all() and (r_on["num_corr"] == r_l2["num_corr"]).all()
    assert_pose_close(np.c_[r_l2["R"], r_l2["t"]], np.c_[r_on["R"], r_on["t"]], "l2 vs on-chip")
    for p in range(0, 24, 6):
        tr = oracle.trace(*b.pair(p), b.init_Rt()[p], max_iter=30, min_delta=-2.0)
        so = b.scene_off
        assert (c_l2[0, so[2 * p]:so[2 * p + 2]] == tr["corr"][-1]).all()
        assert_pose_close(np.r_[r_l2[p]["R"], r_l2[p]["t"]], tr["Rt"], "l2 pair %d" % p)


def test_forced_l2_lockstep_with_ties(l2ctx, oracle):
    """Lattice model: many exact float ties -> the kd-tree replay runs on the global workspace."""
    g = np.arange(12, dtype=np.float64) * 0.5
    model = np.stack(np.meshgrid(g, g), -1).reshape(-1, 2)
    scene = model + 0.25                           # every query equidistant from 4 lattice points
    pair = (model[:40], model[40:], scene[:40], scene[40:])
    Rt0 = np.array([1.0, 0, 0, 1, 0, 0])
    tr = oracle.trace(*pair, Rt0, max_iter=3, min_delta=-2.0)
    corr, res = gpu_lockstep_corr(l2ctx, pair, tr["Rt_trace"][:-1])
    assert (corr == tr["corr"]).all()
    assert (res["status"] & capi.ST_TIE_REPLAY).any()


def test_l2_tie_replay_on_large_lattice(ctx, clctx, oracle):
    """16384-point lattice (L2 variant by size): every query is equidistant from four model points, so each one is
    answered by the cooperative kd-tree replay (first iteration) -- the data-parallel partition (kdreplay.h: partition_ranked) splits
    nodes from 14384 points down to the leaves, the 2000-point class exercises ranges shorter than the CTA.  The
    correspondences of two lock-step iterations must be the reference's, index for index."""
    g = np.arange(128, dtype=np.float64) * 0.1
    model = np.stack(np.meshgrid(g, g), -1).reshape(-1, 2)
    rng = np.random.default_rng(11)
    pick = np.sort(rng.choice(len(model), 1500, replace=False))
    scene = model[pick] + 0.05
    nGA = 2000
    pair = (model[:nGA], model[nGA:], scene[pick < nGA], scene[pick >= nGA])
    Rt0 = np.array([1.0, 0, 0, 1, 0, 0])
    tr = oracle.trace(*pair, Rt0, max_iter=2, min_delta=-2.0)
    for c, variant in ((ctx, 1), (clctx, 4)):
        corr, res = gpu_lockstep_corr(c, pair, tr["Rt_trace"][:-1])
        assert c.launch_info().kernel_variant == variant
        assert (corr == tr["corr"]).all()
        assert (res["status"][0] & capi.ST_TIE_REPLAY) and not (res["status"] & capi.ST_TREE_OVERFLOW).any()
        assert int(res["n_tie_replays"][0]) == int(tr["tie"][0].sum()) >= 900      # (float rounding breaks a third of the ties)


def test_c4_dense_pair(ctx, clctx, oracle):
    """Config C4: 32768-beam pairs (do not fit shared memory -> L2 variant; cluster variant with CCICP_CLUSTER=1),
    30 fixed iterations, with the information matrix."""
    b = scangen.generate(5000, 2, config=4)
    assert b.max_model > 30000
    prm = capi.fixed_iter_params(30, flags=capi.F_INFO)
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    assert ctx.launch_info().kernel_variant == 1
    res1, corr1 = clctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    assert clctx.launch_info().kernel_variant == 4
    assert (corr == corr1).all() and (res["num_corr"] == res1["num_corr"]).all()
    assert_pose_close(np.c_[res["R"], res["t"]], np.c_[res1["R"], res1["t"]], "cluster vs L2")
    np.testing.assert_allclose(res["info"], res1["info"], rtol=1e-9)
    np.testing.assert_allclose(res["score"], res1["score"], rtol=1e-12)
    for p in range(2):
        tr = oracle.trace(*b.pair(p), b.init_Rt()[p], max_iter=30, min_delta=-2.0)
        so = b.scene_off
        got, want = corr[0, so[2 * p]:so[2 * p + 2]], tr["corr"][-1]
        assert (got == want).all(), "pair %d: %d of %d indices differ" % (p, (got != want).sum(), len(want))
        assert res[p]["num_corr"] == tr["ncorr"][-1]
        assert_pose_close(np.r_[res[p]["R"], res[p]["t"]], tr["Rt"], "C4 pair %d" % p)


def test_c4_lockstep_first_iterations(ctx, oracle):
    b = scangen.generate(5100, 1, config=4)
    tr = oracle.trace(*b.pair(0), b.init_Rt()[0], max_iter=4, min_delta=-2.0)
    corr, res = gpu_lockstep_corr(ctx, b.pair(0), tr["Rt_trace"][:-1])
    assert (corr == tr["corr"]).all(), "index mismatch in %d entries" % (corr != tr["corr"]).sum()
    assert (res["num_corr"] == tr["ncorr"]).all()


def test_forced_l2_exact_sums_bit_identical(l2ctx, oracle):
    b = scangen.generate(1000, 6)
    prm = capi.fixed_iter_params(30, flags=capi.F_EXACT_SUMS)
    res = l2ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    for p in range(6):
        r = oracle.fit(*b.pair(p), b.init_Rt()[p])
        assert np.array_equal(np.r_[res[p]["R"], res[p]["t"]], r["Rt"])
