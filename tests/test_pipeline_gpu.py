"""The chunked host pipeline of ccicp_register_batch (H2D of chunk c+1 under the kernel of chunk c, short first chunks,
four compute streams) must return exactly the records of a single launch -- with page-locked buffers (copied in place)
and with pageable ones (init / out staged through the context's pinned memory)."""
import os

import numpy as np
import pytest

from slam_b200 import capi, scangen
from helpers import assert_pose_close

pytestmark = pytest.mark.gpu


@pytest.fixture()
def chunky():
    os.environ["CCICP_CHUNK_PAIRS"] = "64"
    try:
        c = capi.Context(0)
    finally:
        del os.environ["CCICP_CHUNK_PAIRS"]
    yield c
    c.close()


def _same(a, b):
    for f in ("R", "t", "num_corr", "iters", "status", "info"):
        assert np.array_equal(a[f], b[f]), f


@pytest.mark.parametrize("n_pairs", [130, 300, 1100])
def test_chunked_equals_single_launch(ctx, chunky, oracle, n_pairs):
    b = scangen.generate(4200, n_pairs)
    prm = capi.fixed_iter_params(30, flags=capi.F_INFO)
    args = (b.model_xy, b.model_off, b.scene_xy, b.scene_off, b.init_Rt(), prm)
    ref = ctx.register_batch(*args)                       # n_pairs < 2 * 256 (130, 300) -> one launch; 1100 -> chunked too
    nl = ctx.launch_info().n_launches                     # chunk kernels + ONE retry collect + ONE exact-order re-run per call
    assert nl == 3 if n_pairs < 512 else 4 <= nl <= 18
    got = chunky.register_batch(*args)                    # pageable arrays: staging path
    assert chunky.launch_info().n_launches > 1
    _same(ref, got)
    # page-locked arrays: copied in place
    bp = scangen.generate(4200, n_pairs, pinned=capi.pinned_array)
    init = capi.pinned_array((n_pairs, 6), np.float64)
    init[:] = bp.init_Rt()
    got2 = chunky.register_batch(bp.model_xy, bp.model_off, bp.scene_xy, bp.scene_off, init, prm)
    _same(ref, got2)
    for p in (0, n_pairs // 2, n_pairs - 1):
        r = oracle.fit(*b.pair(p), b.init_Rt()[p], max_iter=30, min_delta=-2.0)
        assert got[p]["num_corr"] == r["num_corr"]
        assert_pose_close(np.r_[got[p]["R"], got[p]["t"]], r["Rt"], "pair %d" % p)
