"""Scan-table form of the batch call (ccicp_scans_upload + ccicp_register_scan_pairs): scans are uploaded once and pairs
are (model scan, scene scan) index lists, the shape of graph_slam's loop-closure candidates where every key frame takes
part in K + 1 registrations (graphSlamTools.h:26-36, graphSlamTools.cpp:72-106).  Must reproduce ccicp_register_batch
on the same pairs written out as CSR blocks, record for record."""
import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import csr_from_pairs

pytestmark = pytest.mark.gpu


def _table(n_worlds, seed0=700, config=3):
    b = scangen.generate(seed0, n_worlds, config=config)
    scans, off = [], [0]
    for p in range(n_worlds):
        mg, mn, sg, sn = b.pair(p)
        for ga, nga in ((mg, mn), (sg, sn)):              # scan 2p = the model scan of world p, scan 2p+1 = its scene scan
            scans += [ga, nga]
            off += [off[-1] + len(ga), off[-1] + len(ga) + len(nga)]
    return b, np.concatenate(scans, 0), np.array(off, np.int32)


@pytest.mark.parametrize("n_seeds,flags", [(1, 0), (1, capi.F_INFO), (4, 0), (4, capi.F_BEST_OF_SEEDS)])
def test_scan_pairs_equal_the_csr_call(ctx, n_seeds, flags):
    n_worlds = 10
    b, xy, off = _table(n_worlds)
    rng = np.random.default_rng(1)
    model_scan = np.r_[2 * np.arange(n_worlds), 2 * rng.integers(0, n_worlds, 25)].astype(np.int32)      # reuse: 35 pairs, 20 scans
    scene_scan = (model_scan + 1).astype(np.int32)
    world = model_scan // 2
    P = len(model_scan)
    init = np.repeat(b.init_Rt()[world], n_seeds, 0).reshape(P, n_seeds, 6)
    init[:, :, 4:] += rng.normal(0, 0.05, (P, n_seeds, 2))                                         # every problem its own start
    init = np.ascontiguousarray(init.reshape(-1, 6))
    prm = capi.fixed_iter_params(20, flags=flags)
    ctx.scans_upload(xy, off)
    got = ctx.register_scan_pairs(model_scan, scene_scan, init, prm, n_seeds=n_seeds)
    mxy, moff, sxy, soff = csr_from_pairs([b.pair(int(w)) for w in world])
    want = ctx.register_batch(mxy, moff, sxy, soff, init, prm, n_seeds=n_seeds)
    assert len(got) == len(want)
    for f in ("R", "t", "info", "num_corr", "iters", "status", "best_seed", "score", "last_delta"):
        assert np.array_equal(got[f], want[f]), f
    assert got["num_corr"].min() > 500


def test_scan_table_argument_checks(ctx):
    b, xy, off = _table(2)
    ctx.scans_upload(xy, off)
    prm = capi.fixed_iter_params(3)
    I = np.array([[1.0, 0, 0, 1, 0, 0]])
    with pytest.raises(capi.CcicpError):
        ctx.register_scan_pairs([0], [4], I, prm)                       # scan 4 does not exist (4 scans: 0..3)
    with pytest.raises(capi.CcicpError):
        ctx.register_scan_pairs([-1], [1], I, prm)
    # a scan against itself: zero motion, every point its own neighbour
    r = ctx.register_scan_pairs([0], [0], I, prm)
    assert r[0]["num_corr"] == off[2] - off[0] and np.abs(r[0]["t"]).max() < 1e-12
