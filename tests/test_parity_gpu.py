"""GPU parity tests: CUDA path (through the C ABI) vs the CPU oracle on identical seeded inputs.
Integer work (correspondence indices, inlier sets, counts) must be bit-exact; transforms within
1e-5 m / 1e-6 rad (BASELINE.json north_star)."""
import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import assert_pose_close, csr_from_pairs, gpu_lockstep_corr

pytestmark = pytest.mark.gpu


def _oracle_batch(oracle, batch, idx, **kw):
    out = []
    for p in idx:
        out.append(oracle.trace(*batch.pair(p), batch.init_Rt()[p], **kw))
    return out


@pytest.mark.parametrize("class_rule", [0, 1, 2])
def test_c1_single_pair_free_running(ctx, oracle, class_rule):
    """Config C1: one 1081-beam pair, 30 fixed iterations, identity init."""
    b = scangen.generate(1, 1, params=scangen.default_params(1, class_rule=class_rule))
    tr = oracle.trace(*b.pair(0), b.init_Rt()[0], max_iter=30, min_delta=-2.0, want_info=True)
    prm = capi.fixed_iter_params(30, flags=capi.F_INFO)
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    r = res[0]
    assert r["iters"] == 30 and (int(r["status"]) & ~capi.ST_TIE_REPLAY) == 0
    assert r["num_corr"] == tr["ncorr"][-1]
    assert (corr[0] == tr["corr"][-1]).all()
    assert_pose_close(np.r_[r["R"], r["t"]], tr["Rt"], "C1 rule %d" % class_rule)
    np.testing.assert_allclose(r["info"], tr["info"], rtol=1e-6, atol=1e-6 * np.abs(tr["info"]).max())
    assert r["last_delta"] == pytest.approx(tr["delta"][-1], rel=1e-6, abs=1e-12)
    score = tr["dis"][-1][tr["corr"][-1] >= 0].astype(np.float64).mean()
    assert r["score"] == pytest.approx(score, rel=1e-12)


@pytest.mark.parametrize("seed,config", [(1, 1), (7, 1), (1000003, 3)])
def test_lockstep_bit_exact_every_iteration(ctx, oracle, seed, config):
    """Lock-step: from the oracle's running R,t of EVERY iteration, one GPU fitStep must reproduce the
    reference's per-point correspondence index and inlier mask exactly, and its updated R,t within tolerance."""
    b = scangen.generate(seed, 1, config=config)
    tr = oracle.trace(*b.pair(0), b.init_Rt()[0], max_iter=30, min_delta=-2.0)
    corr, res = gpu_lockstep_corr(ctx, b.pair(0), tr["Rt_trace"][:-1])
    assert corr.shape == tr["corr"].shape
    assert (corr == tr["corr"]).all(), "index mismatch in %d entries" % (corr != tr["corr"]).sum()
    assert (res["num_corr"] == tr["ncorr"]).all()
    assert_pose_close(np.c_[res["R"], res["t"]], tr["Rt_trace"][1:], "lockstep update")


def test_c2_batch_free_running(ctx, oracle):
    """Config C2 shape (odometry pairs), 96-pair sample against the oracle."""
    b = scangen.generate(1000, 96)
    prm = capi.fixed_iter_params(30)
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    trs = _oracle_batch(oracle, b, range(96), max_iter=30, min_delta=-2.0)
    ref_Rt = np.stack([t["Rt"] for t in trs])
    assert_pose_close(np.c_[res["R"], res["t"]], ref_Rt, "C2 batch")
    assert (res["num_corr"] == np.array([t["ncorr"][-1] for t in trs])).all()
    for p, t in enumerate(trs):
        so = b.scene_off
        assert (corr[0, so[2 * p]:so[2 * p + 2]] == t["corr"][-1]).all(), "pair %d" % p


def test_default_params_early_exit(ctx, oracle):
    """Reference defaults (max_iter 20, min_delta 1e-6): same number of fitStep calls, same result."""
    b = scangen.generate(40, 16, params=scangen.default_params(1, off_xy=0.05, off_th=0.01))
    prm = capi.default_params(max_iter=200)
    res = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    for p in range(16):
        tr = oracle.trace(*b.pair(p), b.init_Rt()[p], max_iter=200, min_delta=1e-6)
        assert res[p]["iters"] == tr["iters"], p
        assert_pose_close(np.r_[res[p]["R"], res[p]["t"]], tr["Rt"], "early exit")
        assert bool(res[p]["status"] & capi.ST_CONVERGED) == (tr["iters"] < 200)


def test_c3_loop_closure_sample(ctx, oracle):
    b = scangen.generate(10**6, 32, config=3)
    prm = capi.fixed_iter_params(30)
    res = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    for p in range(32):
        r = oracle.fit(*b.pair(p), b.init_Rt()[p])
        assert_pose_close(np.r_[res[p]["R"], res[p]["t"]], r["Rt"], "C3 pair %d" % p)
        assert res[p]["num_corr"] == r["num_corr"]


@pytest.mark.parametrize("seed,config,n", [(1, 1, 8), (1000, 1, 24), (10**6, 3, 16)])
def test_exact_sums_mode_is_bit_identical(ctx, oracle, seed, config, n):
    """CCICP_F_EXACT_SUMS: sequential-order sums + the reference's NR SVD -> R,t of a free-running 30-iteration fit
    equal the reference's BIT FOR BIT (and with them every correspondence index of every iteration)."""
    b = scangen.generate(seed, n, config=config)
    prm = capi.fixed_iter_params(30, flags=capi.F_EXACT_SUMS)
    res, corr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm, want_corr=True)
    for p in range(n):
        tr = oracle.trace(*b.pair(p), b.init_Rt()[p], max_iter=30, min_delta=-2.0)
        got = np.r_[res[p]["R"], res[p]["t"]]
        assert np.array_equal(got.view(np.uint64), np.asarray(tr["Rt"], np.float64).view(np.uint64)), \
            "pair %d: max |diff| %g" % (p, np.abs(got - tr["Rt"]).max())
        assert res[p]["last_delta"] == tr["delta"][-1]
        so = b.scene_off
        assert (corr[0, so[2 * p]:so[2 * p + 2]] == tr["corr"][-1]).all()
        assert res[p]["num_corr"] == tr["ncorr"][-1]


def test_exact_sums_early_exit_and_lockstep(ctx, oracle):
    """Default stop rule with exact sums: identical iteration count and identical final R,t; one exact fitStep from
    every oracle R,t reproduces the next oracle R,t exactly."""
    b = scangen.generate(40, 8, params=scangen.default_params(1, off_xy=0.05, off_th=0.01))
    prm = capi.default_params(max_iter=200, flags=capi.F_EXACT_SUMS)
    res = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    for p in range(8):
        tr = oracle.trace(*b.pair(p), b.init_Rt()[p], max_iter=200, min_delta=1e-6)
        assert res[p]["iters"] == tr["iters"]
        assert np.array_equal(np.r_[res[p]["R"], res[p]["t"]], tr["Rt"])
    tr = oracle.trace(*b.pair(0), b.init_Rt()[0], max_iter=12, min_delta=-2.0)
    corr, r1 = gpu_lockstep_corr(ctx, b.pair(0), tr["Rt_trace"][:-1], flags=capi.F_EXACT_SUMS)
    assert np.array_equal(np.c_[r1["R"], r1["t"]], tr["Rt_trace"][1:])
