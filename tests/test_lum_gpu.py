"""batched 3-D Lu-Milios edge information (computeEdgeInformationLUM) against the oracle
restatement (parity unpinned: PCL / Eigen are not available to compile the reference function)."""
import numpy as np
import pytest

from oracle import lum_port

pytestmark = pytest.mark.gpu


def cloud_pair(seed, n=1500, noise=0.03):
    rng = np.random.default_rng(seed)
    g = rng.uniform(-20, 20, (n, 2))
    z = 0.3 * np.sin(g[:, 0] * 0.3) + 0.2 * np.cos(g[:, 1] * 0.2)
    tgt = np.c_[g, z].astype(np.float32)
    th = rng.uniform(-0.01, 0.01)
    R = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1]])
    keep = rng.random(n) < 0.8
    src = (tgt[keep] @ R.T + rng.uniform(-0.05, 0.05, 3) + rng.normal(0, noise, (keep.sum(), 3))).astype(np.float32)
    extra = rng.uniform(-40, 40, (n // 10, 3)).astype(np.float32)            # far points: no correspondence
    return np.r_[src, extra], tgt


def test_edge_information_matches_oracle(ctx):
    pairs = [cloud_pair(s, n) for s, n in [(1, 1500), (2, 300), (3, 2500), (4, 2049), (5, 40)]]
    info, num, st = ctx.edge_information_lum([p[0] for p in pairs], [p[1] for p in pairs], 0.75)
    for e, (s, t) in enumerate(pairs):
        wi, wn, ws = lum_port.edge_information(s, t, 0.75)
        assert num[e] == wn and st[e] == ws
        np.testing.assert_allclose(info[e], wi, rtol=2e-5, atol=2e-5 * np.abs(wi).max())
        assert np.array_equal(info[e], info[e].T)


def test_degenerate_edges_give_identity(ctx):
    s, t = cloud_pair(7, 400)
    far = s + np.float32(100.0)                                               # nothing within 0.75 m
    same = t.copy()                                                           # perfect fit: ss == 0
    empty = np.zeros((0, 3), np.float32)
    info, num, st = ctx.edge_information_lum([far, same, empty, s], [t, t, t, empty], 0.75)
    for e in range(4):
        assert st[e] == 1 and np.array_equal(info[e], np.eye(6))
    assert num.tolist() == [0, len(t), 0, 0]
    for e, (a, b) in enumerate([(far, t), (same, t), (empty, t), (s, empty)]):
        wi, wn, ws = lum_port.edge_information(a, b, 0.75)
        assert ws == 1 and wn == num[e]


def test_rejection_rule_restatement():
    assert not lum_port.edge_is_rejected((0, 0, 0.1), (5, -9.9, 0.25))
    assert lum_port.edge_is_rejected((0, 0, 0), (10.5, 0, 0))
    assert lum_port.edge_is_rejected((0, 0, 3.1), (0, 0, -3.0)) is False       # wraps to 0.18 rad
    assert lum_port.edge_is_rejected((0, 0, 3.0), (0, 0, -3.0))                # wraps to 0.28 rad


def test_large_clouds_go_through_the_hash_grid(ctx):
    """Clouds beyond one shared-memory tile (the voxel-filtered Velodyne clouds of graph_slam have tens of thousands of
    points): nearest neighbours come from the 3-D hash grid (lum_grid.cuh) and must give the same information matrix
    as the exhaustive sweep and as the oracle."""
    import os
    from slam_b200 import capi
    pairs = [cloud_pair(11, 20000), cloud_pair(12, 6000), cloud_pair(13, 300)]
    src, tgt = [p[0] for p in pairs], [p[1] for p in pairs]
    info, num, st = ctx.edge_information_lum(src, tgt, 0.75)
    assert ctx.launch_info().kernel_variant == 6 and ctx.launch_info().n_launches == 3
    os.environ["CCICP_LUM_GRID"] = "0"
    try:
        info0, num0, st0 = ctx.edge_information_lum(src, tgt, 0.75)
    finally:
        del os.environ["CCICP_LUM_GRID"]
    assert ctx.launch_info().kernel_variant == 3
    assert np.array_equal(num, num0) and np.array_equal(st, st0) and np.array_equal(info, info0)     # same pairs, same sums
    wi, wn, ws = lum_port.edge_information(pairs[1][0], pairs[1][1], 0.75)
    assert num[1] == wn and st[1] == ws
    np.testing.assert_allclose(info[1], wi, rtol=2e-5, atol=2e-5 * np.abs(wi).max())
    # duplicates and exact ties: lowest target index, like the sweep
    t = np.repeat(cloud_pair(14, 3000)[1], 2, 0)
    s = t[::3] + np.float32(0.01)
    a, na, sa = ctx.edge_information_lum([s], [t], 0.75)
    os.environ["CCICP_LUM_GRID"] = "0"
    try:
        b, nb, sb = ctx.edge_information_lum([s], [t], 0.75)
    finally:
        del os.environ["CCICP_LUM_GRID"]
    assert np.array_equal(a, b) and np.array_equal(na, nb)
