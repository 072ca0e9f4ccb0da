"""ccicp_register_batch_device[_ex]: device pointers, caller-declared capacities.  Covers the cell-table sizing with
max_cells_hint = 0 for class sizes beyond every step of the table rule (2049, 8193 points), and the
CCICP_ST_HINT_TOO_SMALL path for undersized max_cells_hint / max_model / max_scene (no write past the carved
working set, R,t untouched, corr_idx -1)."""
import numpy as np
import pytest

from slam_b200 import capi
from helpers import csr_from_pairs

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _cloud(rng, n, scale=20.0):
    return rng.uniform(-scale, scale, (n, 2))


def _pair(rng, nga, nnga, ns_ga, ns_nga):
    mg, mn = _cloud(rng, nga), _cloud(rng, nnga)
    sg = mg[rng.choice(len(mg), ns_ga)] + rng.normal(0, 0.02, (ns_ga, 2)) if nga else np.zeros((0, 2))
    sn = mn[rng.choice(len(mn), ns_nga)] + rng.normal(0, 0.02, (ns_nga, 2)) if nnga else np.zeros((0, 2))
    return mg, mn, sg, sn


def _device_call(ctx, pairs, prm, max_model, max_scene, hint, want_corr=True):
    dev = torch.device("cuda", 0)
    mxy, moff, sxy, soff = csr_from_pairs(pairs)
    n = len(pairs)
    init = np.tile(np.array([1.0, 0, 0, 1, 0.01, -0.02]), (n, 1))
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
    d = [t(mxy, np.float64), t(moff, np.int32), t(sxy, np.float64), t(soff, np.int32), t(init, np.float64)]
    d_out = torch.zeros(n * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    d_corr = torch.full((int(soff[-1]),), -7, dtype=torch.int32, device=dev)
    ctx.register_batch_device(n, 1, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), d[4].data_ptr(),
                              prm, d_out.data_ptr(), d_corr.data_ptr() if want_corr else 0, max_model, max_scene,
                              None, max_cells_hint=hint)
    ctx.sync()
    torch.cuda.synchronize()
    res = d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    return res, d_corr.cpu().numpy(), (mxy, moff, sxy, soff, init)


@pytest.mark.parametrize("nga,nnga", [(2500, 2500), (2049, 2951), (8200, 8200), (8193, 9000)])
def test_device_api_hint_zero_sizes_the_cell_table(ctx, nga, nnga):
    """max_cells_hint = 0: the library must assume the worst split of max_model over the two classes, including
    both classes above 2048 and above 8192 points (grid_side steps to 128 and 512 cells per side there)."""
    rng = np.random.default_rng(nga)
    pairs = [_pair(rng, nga, nnga, 300, 300)]
    prm = capi.fixed_iter_params(4)
    res, corr, host = _device_call(ctx, pairs, prm, nga + nnga, 600, 0)
    assert not (res["status"] & capi.ST_HINT_TOO_SMALL).any(), "worst-case cell count underestimated"
    want, wcorr = ctx.register_batch(*host, prm, want_corr=True)
    assert (res["num_corr"] == want["num_corr"]).all() and (res["R"] == want["R"]).all() and (res["t"] == want["t"]).all()
    assert (corr == wcorr[0]).all() and res["num_corr"][0] > 500


@pytest.mark.parametrize("what", ["cells", "model", "scene"])
def test_undersized_declarations_are_reported_not_written_through(ctx, what):
    rng = np.random.default_rng(3)
    small = _pair(rng, 40, 200, 30, 150)
    big = _pair(rng, 600, 1500, 400, 1400)        # needs 2 x 64^2 cells... (600 -> 64, 1500 -> 64)
    pairs = [small, big, small]
    prm = capi.fixed_iter_params(5)
    mm, ms, hint = 2100, 1800, 0
    if what == "cells":
        hint = capi.load().ccicp_grid_cells(40) + capi.load().ccicp_grid_cells(200)
    elif what == "model":
        mm = 240
    else:
        ms = 180
    res, corr, host = _device_call(ctx, pairs, prm, mm, ms, hint)
    mxy, moff, sxy, soff, init = host
    assert res[1]["status"] == capi.ST_HINT_TOO_SMALL and res[1]["iters"] == 0 and res[1]["num_corr"] == 0
    assert (np.r_[res[1]["R"], res[1]["t"]] == init[1]).all()
    assert (corr[soff[2]:soff[4]] == -1).all()
    want, wcorr = ctx.register_batch(*host, prm, want_corr=True)
    for p in (0, 2):                              # the neighbours are registered normally
        assert res[p]["status"] == want[p]["status"] and res[p]["num_corr"] == want[p]["num_corr"]
        assert (res[p]["R"] == want[p]["R"]).all()
        assert (corr[soff[2 * p]:soff[2 * p + 2]] == wcorr[0, soff[2 * p]:soff[2 * p + 2]]).all()


def test_zero_iterations_reports_nothing(ctx):
    """max_iter <= 0: no fitStep runs; the reference's fitIterate does nothing (icp.cpp:116-122) -- no NO_CORR flag."""
    rng = np.random.default_rng(5)
    pairs = [_pair(rng, 40, 200, 30, 150)]
    mxy, moff, sxy, soff = csr_from_pairs(pairs)
    Rt0 = np.array([[0.8, -0.6, 0.6, 0.8, 1.0, 2.0]])
    res = ctx.register_batch(mxy, moff, sxy, soff, Rt0, capi.default_params(max_iter=0))
    assert res[0]["status"] == 0 and res[0]["iters"] == 0 and (np.r_[res[0]["R"], res[0]["t"]] == Rt0[0]).all()
