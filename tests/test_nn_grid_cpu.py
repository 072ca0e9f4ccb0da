"""The product's grid search (slam_b200/csrc/nn_grid.h: nn_search, both instantiations) compiled for the host and run on
the CPU against a brute-force float arg-min: every query whose true nearest neighbour is inside the gate must get the
exact float minimum, the right slot (or a tie flag when the minimum is shared) -- for cold queries and for queries
warm-started from arbitrary slots.  No GPU needed; the kernels include the same header."""
import ctypes as C
import os

import numpy as np
import pytest

from slam_b200 import scangen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBP = os.path.join(ROOT, "tests", "host", "libnn_grid_host.so")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIBP):
        pytest.fail("tests/host/libnn_grid_host.so missing: run python -c 'import __graft_entry__ as g; g.build()'")
    return C.CDLL(LIBP)


def P(a):
    return a.ctypes.data_as(C.c_void_p)


def run(lib, mGA, mNGA, q, cls, warm, gate, dense):
    xy = np.ascontiguousarray(np.concatenate([mGA.reshape(-1, 2), mNGA.reshape(-1, 2)], 0), np.float64)
    n = len(xy)
    mpts = np.zeros((max(n, 1), 2), np.float32); morig = np.zeros(max(n, 1), np.uint16)
    cstart = np.zeros(2 * 1024 * 1024 + 8, np.uint16); grid = np.zeros(2 * lib.nng_gridp_size(), np.uint8)
    lib.nng_build(P(xy), len(mGA), len(mNGA), P(mpts), P(morig), P(cstart), P(grid))
    q = np.ascontiguousarray(q, np.float32); nq = len(q)
    qx, qy = np.ascontiguousarray(q[:, 0]), np.ascontiguousarray(q[:, 1])
    cls = np.ascontiguousarray(cls, np.uint8)
    first = np.where(cls == 1, len(mGA) if len(mGA) > 3 else 0, 0)      # slots: a skipped class occupies none
    cnt = np.where(cls == 1, len(mNGA), len(mGA))
    w = np.where((warm >= 0) & (cnt > 3), first + np.clip(warm, 0, np.maximum(cnt - 1, 0)), 0xFFFF).astype(np.uint16)
    slot = np.zeros(nq, np.int32); dist = np.zeros(nq, np.float32); tie = np.zeros(nq, np.uint8); ev = np.zeros(nq, np.int32)
    lib.nng_query(P(mpts), P(cstart), P(grid), int(dense), nq, P(cls), P(qx), P(qy), P(w), C.c_float(gate), P(slot), P(dist), P(tie), P(ev))
    bs = np.zeros(nq, np.int32); bd = np.zeros(nq, np.float32); sh = np.zeros(nq, np.int32)
    lib.nng_brute(P(mpts), P(cstart), P(grid), nq, P(cls), P(qx), P(qy), P(bs), P(bd), P(sh))
    return slot, dist, tie, ev, bs, bd, sh


def check(lib, mGA, mNGA, q, cls, gate=5.0, seed=0):
    rng = np.random.default_rng(seed)
    gate_f = np.float32(gate)
    for dense in (0, 1):
        for warm in (np.full(len(q), -1), rng.integers(0, 1 << 15, len(q))):
            slot, dist, tie, ev, bs, bd, sh = run(lib, mGA, mNGA, q, cls, warm, gate_f, dense)
            inside = (bs >= 0) & (bd < gate_f)
            assert (dist[inside] == bd[inside]).all(), "float minimum differs"
            uniq = inside & (sh == 1)
            assert (slot[uniq] == bs[uniq]).all() and (tie[uniq] == 0).all()
            shared = inside & (sh > 1)
            assert (tie[shared] == 1).all(), "a shared minimum must be flagged for the kd-tree replay"
            found = slot >= 0
            assert (dist[found] >= bd[found]).all()


def test_scan_pairs_cold_and_warm(lib):
    b = scangen.generate(1000, 3)
    for p in range(3):
        mg, mn, sg, sn = b.pair(p)
        scene = np.concatenate([sg, sn], 0)
        cls = (np.arange(len(scene)) >= len(sg)).astype(np.uint8)
        for shift in (0.0, 0.4, 3.0):
            check(lib, mg, mn, scene + shift, cls, seed=p)


@pytest.mark.parametrize("n", [5, 40, 200, 700, 2500])
def test_random_clustered_and_lattice_clouds(lib, n):
    rng = np.random.default_rng(n)
    uni = rng.uniform(-20, 20, (n, 2))
    clu = rng.normal(0, 0.3, (n, 2)) + rng.integers(-3, 4, (n, 1)) * 5.0
    lat = np.round(rng.uniform(-6, 6, (n, 2)) * 2) / 2            # duplicates + exact ties
    for k, m in enumerate((uni, clu, lat)):
        nga = n // 3
        q = np.r_[m[rng.integers(0, n, 300)] + rng.normal(0, 0.2, (300, 2)), rng.uniform(-25, 25, (200, 2)),
                  np.round(rng.uniform(-6, 6, (200, 2)) * 4) / 4]
        cls = rng.integers(0, 2, len(q)).astype(np.uint8)
        for gate in (5.0, 0.05, 400.0):
            check(lib, m[:nga], m[nga:], q, cls, gate=gate, seed=k)


def test_degenerate_extents(lib):
    line = np.c_[np.linspace(-5, 5, 300), np.zeros(300)]           # zero extent in y
    point = np.tile([[1.5, -2.5]], (50, 1))                        # all points identical
    rng = np.random.default_rng(3)
    q = rng.uniform(-6, 6, (400, 2)); cls = rng.integers(0, 2, 400).astype(np.uint8)
    check(lib, line[:100], line[100:], q, cls)
    check(lib, point[:10], point[10:], q, cls, gate=100.0)
