"""ccicp_register_batch_device_static: the allocation-free device-pointer call (scratch in one file-scope device array of
CCICP_STATIC_SCRATCH_BYTES, no cudaMalloc inside the call).  It has no counterpart in the reference; its records must be
bit-identical to ccicp_register_batch_device's on the same inputs -- on two calls in a row, the second one finding the
first one's counters, problem list and kd-tree nodes in the shared array.  Collected last on
purpose."""
import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import csr_from_pairs

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

FIELDS = ("R", "t", "num_corr", "iters", "status", "score", "info", "last_delta", "n_tie_replays")


def _to_dev(a, dt):
    return torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1).copy()).to(torch.device("cuda", 0))


def _both_calls(ctx, mxy, moff, sxy, soff, init, prm, max_model, max_scene, want_corr=True):
    """Runs the existing entry point, then the static one, on the same device buffers; returns (want, got, corr_want, corr_got)."""
    n = len(moff) // 2
    d = [_to_dev(mxy, np.float64), _to_dev(moff, np.int32), _to_dev(sxy, np.float64), _to_dev(soff, np.int32),
         _to_dev(init, np.float64)]
    nb = max(n, 1) * capi.RESULT_DTYPE.itemsize
    outs, corrs = [], []
    for static in (False, True):
        d_out = torch.full((nb,), 0xEE, dtype=torch.uint8, device="cuda:0")
        d_corr = torch.full((max(int(soff[-1]), 1),), -7, dtype=torch.int32, device="cuda:0")
        torch.cuda.synchronize()
        ptrs = [t.data_ptr() for t in d]
        if static:
            ctx.register_batch_device_static(n, *ptrs, prm, d_out.data_ptr(), d_corr.data_ptr() if want_corr else 0,
                                             max_model, max_scene)
        else:
            ctx.register_batch_device(n, 1, *ptrs, prm, d_out.data_ptr(), d_corr.data_ptr() if want_corr else 0,
                                      max_model, max_scene)
        ctx.sync()
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy()[:n * capi.RESULT_DTYPE.itemsize].view(capi.RESULT_DTYPE).copy())
        corrs.append(d_corr.cpu().numpy()[:int(soff[-1])].copy())
    return outs[0], outs[1], corrs[0], corrs[1]


def _assert_same(want, got, cw, cg, what):
    for f in FIELDS:
        a, b = want[f].reshape(len(want), -1), got[f].reshape(len(got), -1)
        same = (a == b) | ((a != a) & (b != b)) if a.dtype.kind == "f" else (a == b)      # bit-exact; NaN equals NaN
        bad = np.flatnonzero(~same.all(axis=1))
        assert len(bad) == 0, "%s: field %s differs from ccicp_register_batch_device in records %s" % (what, f, bad[:8])
    assert (cw == cg).all(), "%s: correspondence indices differ" % what


def _scan_batch(seed, n_pairs):
    b = scangen.generate(seed, n_pairs)
    max_model = int(np.diff(b.model_off[::2]).max())
    max_scene = int(np.diff(b.scene_off[::2]).max())
    return (b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt()), max_model, max_scene


def _degenerate_pairs(rng, n):
    """Pairs whose gate keeps one or two correspondences: H is rank deficient, the record comes from the exact-order
    re-run (its problem list lives in the static array as well)."""
    pairs = []
    for _ in range(n):
        mg = rng.uniform(-20, 20, (40, 2)); mn = rng.uniform(-20, 20, (60, 2))
        sg = np.r_[mg[:1] + 0.01, rng.uniform(200, 300, (4, 2))]          # one GA point inside the gate, the rest far away
        sn = rng.uniform(200, 300, (3, 2))
        pairs.append((mg, mn, sg, sn))
    return pairs


def test_static_call_equals_device_call_twice_in_a_row(ctx):
    prm = capi.fixed_iter_params(30, flags=capi.F_INFO)
    # call 1: 300 odometry scan pairs (1081 beams): more pairs than CTAs, so the work counter in the array is exercised
    host1, mm1, ms1 = _scan_batch(1000, 300)
    w1, g1, cw1, cg1 = _both_calls(ctx, *host1, prm, mm1, ms1)
    _assert_same(w1, g1, cw1, cg1, "first call")
    assert (g1["num_corr"] > 300).all() and (g1["iters"] == 30).all()
    # call 2, right behind it: other scans, fewer pairs, early exit allowed -- stale contents of the array must not matter
    host2, mm2, ms2 = _scan_batch(77000, 41)
    prm2 = capi.default_params(max_iter=40)
    w2, g2, cw2, cg2 = _both_calls(ctx, *host2, prm2, mm2, ms2)
    _assert_same(w2, g2, cw2, cg2, "second call")
    # call 3: the first batch again -> the same records as call 1
    _, g3, _, cg3 = _both_calls(ctx, *host1, prm, mm1, ms1)
    _assert_same(g1, g3, cg1, cg3, "repeat of the first call")


def test_static_call_runs_the_exact_order_rerun_from_the_array(ctx):
    rng = np.random.default_rng(11)
    scans, mm, ms = _scan_batch(4242, 24)
    b = scangen.generate(4242, 24)
    pairs = [b.pair(p) for p in range(24)]
    deg = _degenerate_pairs(rng, 9)
    mixed = pairs[:10] + deg[:5] + pairs[10:] + deg[5:]
    mxy, moff, sxy, soff = csr_from_pairs(mixed)
    init = np.tile(np.array([1.0, 0, 0, 1, 0.0, 0.0]), (len(mixed), 1))
    prm = capi.fixed_iter_params(6)
    for rep in range(2):
        w, g, cw, cg = _both_calls(ctx, mxy, moff, sxy, soff, init, prm, max(mm, 100), max(ms, 8))
        _assert_same(w, g, cw, cg, "mixed batch, call %d" % rep)
        assert (g["status"] & (capi.ST_EXACT_RETRY | capi.ST_DEGENERATE)).astype(bool).sum() >= 9, "no problem took the re-run path"


def test_static_call_rejects_sizes_beyond_its_limits(ctx):
    prm = capi.fixed_iter_params(2)
    host, mm, ms = _scan_batch(5, 2)
    d = [_to_dev(host[0], np.float64), _to_dev(host[1], np.int32), _to_dev(host[2], np.float64), _to_dev(host[3], np.int32),
         _to_dev(host[4], np.float64)]
    d_out = torch.zeros(2 * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda:0")
    ptrs = [t.data_ptr() for t in d]
    for kw, word in ((dict(max_model=capi.STATIC_MAX_POINTS + 1, max_scene=ms), "CCICP_STATIC_MAX_POINTS"),
                     (dict(max_model=mm, max_scene=capi.STATIC_MAX_POINTS + 1), "CCICP_STATIC_MAX_POINTS")):
        with pytest.raises(capi.CcicpError) as e:
            ctx.register_batch_device_static(2, *ptrs, prm, d_out.data_ptr(), 0, kw["max_model"], kw["max_scene"])
        assert e.value.code == capi.E_TOO_LARGE and word in str(e.value)
    with pytest.raises(capi.CcicpError) as e:
        ctx.register_batch_device_static(capi.STATIC_MAX_PAIRS + 1, *ptrs, prm, d_out.data_ptr(), 0, mm, ms)
    assert e.value.code == capi.E_TOO_LARGE and "CCICP_STATIC_MAX_PAIRS" in str(e.value)
    with pytest.raises(capi.CcicpError) as e:
        ctx.register_batch_device_static(2, *ptrs, capi.fixed_iter_params(2, flags=capi.F_EXACT_SUMS), d_out.data_ptr(), 0, mm, ms)
    assert e.value.code == capi.E_INVALID
    # and the call still works afterwards
    ctx.register_batch_device_static(2, *ptrs, prm, d_out.data_ptr(), 0, mm, ms)
    ctx.sync()
    res = d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    assert (res["iters"] == 2).all()
