"""GPU parity on the edge cases the reference's control flow distinguishes (SURVEY.md 8(a) a1,a3,a5,a6)."""
import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import assert_pose_close, csr_from_pairs, gpu_lockstep_corr

pytestmark = pytest.mark.gpu
I6 = np.array([1.0, 0, 0, 1, 0, 0])
E = np.zeros((0, 2))


def _run(ctx, pairs, Rt0, prm, n_seeds=1, want_corr=True):
    mxy, moff, sxy, soff = csr_from_pairs(pairs)
    return ctx.register_batch(mxy, moff, sxy, soff, Rt0, prm, n_seeds=n_seeds, want_corr=want_corr)


def test_ties_follow_reference_traversal(ctx, oracle):
    """Lattice model with duplicate points, queries on lattice midpoints: the float minimum is shared by 2-4
    model points for most queries, so the answer is decided by the reference's kd-tree traversal order."""
    rng = np.random.default_rng(5)
    model = np.round(rng.uniform(-12, 12, (900, 2)))
    scene = np.round(rng.uniform(-12, 12, (700, 2)) * 2) / 2
    pair = (model[:300], model[300:], scene[:250], scene[250:])
    tr = oracle.trace(*pair, I6, max_iter=1, min_delta=-2.0)
    assert tr["tie"].sum() > 100
    corr, res = gpu_lockstep_corr(ctx, pair, I6[None])
    assert (corr == tr["corr"]).all(), "%d tie mismatches" % (corr != tr["corr"]).sum()
    assert res[0]["status"] & capi.ST_TIE_REPLAY and res[0]["n_tie_replays"] >= tr["tie"][0][tr["corr"][0] >= 0].sum()
    # and the whole iteration sequence
    tr = oracle.trace(*pair, I6, max_iter=10, min_delta=-2.0)
    res, corr = _run(ctx, [pair], I6[None], capi.fixed_iter_params(10))
    assert (corr[0] == tr["corr"][-1]).all()
    assert_pose_close(np.r_[res[0]["R"], res[0]["t"]], tr["Rt"], "ties")


def test_too_few_points_leave_pose_untouched(ctx):
    rng = np.random.default_rng(0)
    big = rng.uniform(-5, 5, (50, 2))
    Rt0 = np.array([0.8, -0.6, 0.6, 0.8, 1.5, -2.5])
    pairs = [(big[:2], big[2:4], big[:20], big[20:]),      # 4 model points  (icp.cpp:37-43)
             (big[:20], big[20:], big[:1], big[1:4]),      # 4 scene points  (icp.cpp:100-103)
             (E, E, E, E)]
    res, corr = _run(ctx, pairs, np.tile(Rt0, (3, 1)), capi.fixed_iter_params(5))
    assert res[0]["status"] == capi.ST_TOO_FEW_MODEL and res[1]["status"] == capi.ST_TOO_FEW_SCENE
    assert res[2]["status"] == capi.ST_TOO_FEW_MODEL
    for r in res:
        assert (np.r_[r["R"], r["t"]] == Rt0).all() and r["iters"] == 0 and r["num_corr"] == 0
    assert (corr == -1).all()


def test_small_class_is_skipped(ctx, oracle):
    """A class whose MODEL has <= 3 points is never searched (icpPointToPoint.cpp:59,93)."""
    b = scangen.generate(3, 1)
    mg, mn, sg, sn = b.pair(0)
    pair = (mg[:3], mn, sg, sn)
    tr = oracle.trace(*pair, I6, max_iter=8, min_delta=-2.0)
    res, corr = _run(ctx, [pair], I6[None], capi.fixed_iter_params(8))
    assert (corr[0][:len(sg)] == -1).all() and (corr[0] == tr["corr"][-1]).all()
    assert_pose_close(np.r_[res[0]["R"], res[0]["t"]], tr["Rt"], "skipped class")
    # empty scene class, empty model class
    for pair in [(mg, mn, E, sn), (E, mn, sg, sn), (mg, E, sg, sn)]:
        tr = oracle.trace(*pair, I6, max_iter=5, min_delta=-2.0)
        res, corr = _run(ctx, [pair], I6[None], capi.fixed_iter_params(5))
        if tr["iters"] and len(tr["corr"][-1]):
            assert (corr[0] == tr["corr"][-1]).all()
        if not (tr["status"] & 4):
            assert_pose_close(np.r_[res[0]["R"], res[0]["t"]], tr["Rt"], "empty class")


def test_no_correspondences(ctx, oracle):
    """Everything beyond the gate: fitStep returns -1, R,t untouched; default min_delta stops after 1 call,
    fixed-iteration mode (min_delta < -1) keeps calling (icp.cpp:116-122)."""
    rng = np.random.default_rng(1)
    model = rng.uniform(-5, 5, (200, 2))
    scene = rng.uniform(-5, 5, (150, 2)) + 100.0
    pair = (model[:50], model[50:], scene[:40], scene[40:])
    res, corr = _run(ctx, [pair, pair], np.tile(I6, (2, 1)), capi.default_params(max_iter=7))
    assert res[0]["status"] == capi.ST_NO_CORR and res[0]["iters"] == 1 and res[0]["last_delta"] == -1.0
    assert (np.r_[res[0]["R"], res[0]["t"]] == I6).all() and (corr == -1).all()
    assert np.isinf(res[0]["score"])
    res, _ = _run(ctx, [pair], I6[None], capi.fixed_iter_params(7))
    assert res[0]["iters"] == 7 and res[0]["status"] == capi.ST_NO_CORR


def test_gate_is_strict_on_squared_float_distance(ctx, oracle):
    """keep iff (double)dis < indist with dis the SQUARED float distance (icpPointToPoint.cpp:76)."""
    b = scangen.generate(11, 1, params=scangen.default_params(1, off_xy=1.5, off_th=0.2))
    pair = b.pair(0)
    for indist in (0.01, 0.25, 1.0):
        tr = oracle.trace(*pair, I6, max_iter=6, min_delta=-2.0, indist=indist)
        corr, res = gpu_lockstep_corr(ctx, pair, tr["Rt_trace"][:-1], indist=indist)
        assert (corr == tr["corr"]).all(), indist
    # exactly-on-the-gate point: dis == indist must be rejected, next float below accepted
    model = np.array([[0, 0], [10, 0], [0, 10], [10, 10], [5, 5.0]])
    scene = np.array([[2.0, 0], [10, 0.5], [0, 10.5], [10, 10.5], [5, 5.5]])
    pair = (E, model, E, scene)
    for indist, expect in ((4.0, -1), (np.nextafter(np.float32(4.0), np.float32(5)).astype(np.float64), 0)):
        tr = oracle.trace(*pair, I6, max_iter=1, min_delta=-2.0, indist=float(indist))
        corr, _ = gpu_lockstep_corr(ctx, pair, I6[None], indist=float(indist))
        assert corr[0][0] == expect == tr["corr"][0][0]


def test_reflection_matches_reference(ctx, oracle):
    """det(H) < 0: the reference applies the improper V*U^T (no det fix, icpPointToPoint.cpp:162)."""
    rng = np.random.default_rng(2)
    model = np.c_[np.arange(120) * 0.5 - 30, rng.uniform(-0.1, 0.1, 120)]
    scene = model * np.array([1.0, -1.0]) + rng.normal(0, 1e-3, model.shape)   # mirrored copy, NN = own original
    pair = (E, model, E, scene)
    tr = oracle.trace(*pair, I6, max_iter=3, min_delta=-2.0, indist=100.0)
    assert np.linalg.det(tr["Rt"][:4].reshape(2, 2)) < 0
    res, corr = _run(ctx, [pair], I6[None], capi.default_params(max_iter=3, min_delta=-2.0, indist=100.0))
    assert (corr[0] == tr["corr"][-1]).all()
    assert res[0]["status"] & capi.ST_REFLECTION
    np.testing.assert_allclose(np.r_[res[0]["R"], res[0]["t"]], tr["Rt"], atol=1e-9)


def test_multi_seed_best_of(ctx, oracle):
    """Config C5 shape: hypotheses per pair, arg-min of the mean gated squared distance on device."""
    b = scangen.generate(2000, 4)
    n_seeds = 24
    hyp = np.stack([scangen.hypotheses(int(s), n_seeds) for s in b.seeds])
    hyp[:, 1:, :2] *= 0.08            # keep some hypotheses inside the basin so several seeds qualify
    hyp[:, 1:, 2] = (hyp[:, 1:, 2] - np.pi) * 0.05
    Rt0 = scangen.pose_to_Rt(hyp.reshape(-1, 3))
    prm = capi.fixed_iter_params(12)
    allr = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, Rt0, prm, n_seeds=n_seeds)
    prm_b = capi.fixed_iter_params(12, flags=capi.F_BEST_OF_SEEDS)
    best = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, Rt0, prm_b, n_seeds=n_seeds)
    assert len(best) == 4 and len(allr) == 4 * n_seeds
    for p in range(4):
        ns = len(b.pair(p)[2]) + len(b.pair(p)[3])
        scores = []
        for k in range(n_seeds):
            tr = oracle.trace(*b.pair(p), Rt0[p * n_seeds + k], max_iter=12, min_delta=-2.0)
            g = allr[p * n_seeds + k]
            assert g["num_corr"] == (tr["ncorr"][-1] if tr["iters"] else 0)
            sc = np.inf
            if tr["iters"] and tr["ncorr"][-1] > 0:
                sc = tr["dis"][-1][tr["corr"][-1] >= 0].astype(np.float64).mean()
                assert g["score"] == pytest.approx(sc, rel=1e-9)
                assert_pose_close(np.r_[g["R"], g["t"]], tr["Rt"], "seed %d" % k)
            scores.append(sc if tr["iters"] and tr["ncorr"][-1] >= 0.5 * ns else np.inf)
        assert best[p]["best_seed"] == int(np.argmin(scores))
        k = best[p]["best_seed"]
        assert (best[p]["R"] == allr[p * n_seeds + k]["R"]).all() and best[p]["score"] == allr[p * n_seeds + k]["score"]


def test_multi_seed_first_below_threshold(ctx):
    """global_matching's acceptance rule (global_match.cpp:131-151) -- the FIRST seed whose
    normalised score (score / n_scene) is below the threshold wins, not the arg-min; none below -> arg-min + flag."""
    b = scangen.generate(2000, 6)
    n_seeds = 16
    hyp = np.stack([scangen.hypotheses(int(s), n_seeds) for s in b.seeds])
    hyp[:, 1:, :2] *= 0.05
    hyp[:, 1:, 2] = (hyp[:, 1:, 2] - np.pi) * 0.03
    Rt0 = scangen.pose_to_Rt(hyp.reshape(-1, 3))
    args = (b.model_xy, b.model_off, b.scene_xy, b.scene_off, Rt0)
    allr = ctx.register_batch(*args, capi.fixed_iter_params(12), n_seeds=n_seeds).reshape(6, n_seeds)
    ns = np.diff(b.scene_off)[0::2] + np.diff(b.scene_off)[1::2]
    ok = allr["num_corr"] >= 0.5 * ns[:, None]
    norm = np.where(ok, allr["score"], np.inf) / ns[:, None]
    thr = float(np.median(norm[np.isfinite(norm)]))             # about half of the seeds pass
    prm = capi.fixed_iter_params(12, flags=capi.F_BEST_OF_SEEDS | capi.F_FIRST_BELOW)
    prm.score_threshold = thr
    got = ctx.register_batch(*args, prm, n_seeds=n_seeds)
    for p in range(6):
        below = np.nonzero(norm[p] < thr)[0]
        if len(below):
            assert got[p]["best_seed"] == below[0] and not (got[p]["status"] & capi.ST_NO_SEED_BELOW)
        else:
            assert got[p]["best_seed"] == int(np.argmin(norm[p])) and (got[p]["status"] & capi.ST_NO_SEED_BELOW)
        assert (got[p]["R"] == allr[p, got[p]["best_seed"]]["R"]).all()
    prm.score_threshold = 1e-12                                  # nobody passes: arg-min + flag everywhere
    got = ctx.register_batch(*args, prm, n_seeds=n_seeds)
    assert ((got["status"] & capi.ST_NO_SEED_BELOW) != 0).all()
    assert (got["best_seed"] == np.argmin(np.where(ok, allr["score"], np.inf), 1)).all()
