"""GPU classifyPoints (bin + stable radix sort + neighbour occupancy) against the oracle
restatement: identical point order, flags and source indices (bit-exact: integer / byte work)."""
import numpy as np
import pytest

from oracle import classify_port as cp
from test_classify_cpu import synthetic_cloud

pytestmark = pytest.mark.gpu


def _check(ctx, clouds):
    off = np.cumsum([0] + [len(c) for c in clouds]).astype(np.int32)
    xyz = np.concatenate(clouds, 0) if clouds else np.zeros((0, 3), np.float32)
    got = ctx.classify_points(xyz, off)
    assert len(got) == len(clouds)
    for c, (gx, gg, gs) in zip(clouds, got):
        wx, wg, ws = cp.classify(c)
        assert len(gx) == len(wx)
        assert np.array_equal(gs, ws), "order differs"
        assert np.array_equal(gx.view(np.uint32), wx.view(np.uint32)) and np.array_equal(gg, wg)


def test_single_cloud(ctx):
    _check(ctx, [synthetic_cloud(0, 70000)])


def test_batch_of_ragged_clouds(ctx):
    sizes = [1, 0, 5000, 257, 4096, 4097, 12345, 0, 3]
    clouds = [synthetic_cloud(10 + k, n) if n > 8 else synthetic_cloud(10 + k, 16)[:n] for k, n in enumerate(sizes)]
    _check(ctx, clouds)


def test_edges_nan_and_duplicates(ctx):
    e = np.array([[-300, 0, 0], [299.5, 0, 0], [300, 0, 0], [-299.5, 0.2, 0], [np.nan, 0, 0], [0, np.inf, 0],
                  [1.1, 5.1, 1], [-2.0, 0.1, 2], [1.2, 5.2, 3], [1.1, -5.0, 4], [1.1, 5.1, 5]], np.float32)
    _check(ctx, [e, e[::-1].copy(), np.repeat(e[6:7], 5000, 0)])


def test_many_small_clouds(ctx):
    _check(ctx, [synthetic_cloud(100 + k, 300) for k in range(300)])
