"""GPU classifyPoints (bin + stable radix sort + neighbour occupancy) against the REFERENCE's own
CCICP::classifyPoints (oracle/_ref/libclassify_ref.so: icpTools.cpp:36-103 compiled unmodified, oracle/Makefile) for
point order, coordinates and flags, and against the numpy port (pinned to the same library in
tests/test_classify_cpu.py) for the source indices the reference does not carry.  Bit-exact: integer / byte work."""
import numpy as np
import pytest

from oracle import classify_port as cp
from oracle import refbind
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
        if refbind.classify_available() and len(c) and np.isfinite(c[:, :2]).all():     # the reference itself
            rx, rg = refbind.classify_ref(c)
            assert len(rx) == len(gx)
            assert np.array_equal(gx.view(np.uint32), rx.view(np.uint32)) and np.array_equal(gg, rg)


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


def _device_call(ctx, clouds, misalign_floats=0):
    import torch
    dev = torch.device("cuda", 0)
    off = np.cumsum([0] + [len(c) for c in clouds]).astype(np.int32)
    xyz = np.concatenate(clouds, 0)
    n = len(xyz)
    buf = torch.zeros(3 * n + 4, dtype=torch.float32, device=dev)
    d_xyz = buf[misalign_floats:misalign_floats + 3 * n]
    d_xyz.copy_(torch.from_numpy(xyz.reshape(-1)).to(dev))
    d_off = torch.from_numpy(off).to(dev)
    o_xyz = torch.empty(3 * n, dtype=torch.float32, device=dev); o_ga = torch.empty(n, dtype=torch.uint8, device=dev)
    o_src = torch.empty(n, dtype=torch.int32, device=dev); o_cnt = torch.empty(len(clouds), dtype=torch.int32, device=dev)
    ctx.classify_points_device(len(clouds), d_xyz.data_ptr(), d_off.data_ptr(), n, o_xyz.data_ptr(), o_ga.data_ptr(),
                               o_src.data_ptr(), o_cnt.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    o_xyz, o_ga, o_src, o_cnt = (t.cpu().numpy() for t in (o_xyz.view(n, 3), o_ga, o_src, o_cnt))
    for k, c in enumerate(clouds):
        wx, wg, ws = cp.classify(c)
        assert o_cnt[k] == len(wx)
        sl = slice(off[k], off[k] + len(wx))
        assert np.array_equal(o_src[sl], ws) and np.array_equal(o_ga[sl], wg)
        assert np.array_equal(o_xyz[sl].view(np.uint32), wx.view(np.uint32))


def test_device_api_gather_paths(ctx):
    """Device-pointer entry point, with an 8-byte aligned and a merely 4-byte aligned xyz pointer."""
    clouds = [synthetic_cloud(7, 9001), synthetic_cloud(8, 4096), synthetic_cloud(9, 777)]
    _device_call(ctx, clouds, 0)
    _device_call(ctx, clouds, 1)


def test_one_large_cloud_spans_several_tile_groups(ctx):
    """> 32 tiles: the key kernel splits the cloud over several CTAs that merge their bitmaps with atomics."""
    _device_call(ctx, [synthetic_cloud(11, 300000), synthetic_cloud(12, 50)], 0)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_fuzz_ragged_batches(ctx, seed):
    """Random cloud counts and sizes around the tile (4096) and tile-group (32 tiles) boundaries, clustered points (many
    per bin), points outside the grid and NaNs."""
    rng = np.random.default_rng(1234 + seed)
    sizes = [int(s) for s in rng.choice([0, 1, 2, 31, 4095, 4096, 4097, 8192, 10000, 131072, 131073, 20000], size=rng.integers(3, 9))]
    clouds = []
    for k, n in enumerate(sizes):
        c = synthetic_cloud(1000 * seed + k, max(n, 16))[:n]
        if n > 100:
            c[rng.integers(0, n, n // 50), 0] = np.nan                   # dropped by the bin test
            c[rng.integers(0, n, n // 20)] = c[rng.integers(0, n)]       # duplicates: long runs inside one bin
            c[rng.integers(0, n, n // 50), :2] *= 100.0                   # far outside the 600 m grid
        clouds.append(np.ascontiguousarray(c))
    _device_call(ctx, clouds, 0)
