"""CPU check of the product's tie-replay code (slam_b200/csrc/kdreplay.h compiled for the host): the eager build
must reproduce the reference kd-tree node for node, and both the eager and the lazy traversal must return the
reference's n_nearest(q,1) answer, including on tie-heavy lattice clouds."""
import ctypes as C
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "host", "libkdreplay_host.so")
P = C.c_void_p


@pytest.fixture(scope="module")
def kd():
    if not os.path.exists(LIB):
        pytest.fail("tests/host/libkdreplay_host.so missing: run __graft_entry__.build()")
    return C.CDLL(LIB)


def clouds():
    rng = np.random.default_rng(0)
    for trial in range(24):
        n = int(rng.integers(5, 2500))
        kind = trial % 4
        if kind == 0:
            pts = rng.uniform(-10, 10, (n, 2))
        elif kind == 1:
            pts = np.round(rng.uniform(-40, 40, (n, 2)))          # lattice: duplicates and ties
        elif kind == 2:
            a = np.sort(rng.uniform(-2, 2, n)); r = 5 + 0.01 * rng.standard_normal(n)
            pts = np.stack([r * np.cos(a), r * np.sin(a)], 1)      # a scan-like arc
        else:
            pts = np.round(rng.uniform(-20, 20, (n, 2)) * 2) / 2
        nq = [3, 40, 2500][trial % 3]
        q = np.round(rng.uniform(-42, 42, (nq, 2)) * 2) / 2 if kind in (1, 3) else rng.uniform(-12, 12, (nq, 2))
        yield np.ascontiguousarray(pts, np.float64), np.ascontiguousarray(q, np.float32)


def test_eager_tree_matches_reference(kd, oracle):
    for pts, _ in clouds():
        n = len(pts)
        ind = np.zeros(n, np.int32); nodes = np.zeros(10 * (2 * n + 2), np.float32)
        nn = kd.kdr_tree(pts.ctypes.data_as(P), n, ind.ctypes.data_as(P), nodes.ctypes.data_as(P), 2 * n + 2)
        rn, rind, rnodes = oracle.tree(pts)
        assert nn == rn and (ind == rind).all()
        assert (nodes[:10 * nn].reshape(-1, 10) == rnodes).all()     # == : -0.0 equals +0.0


def test_eager_and_lazy_search_match_reference(kd, oracle):
    for pts, q in clouds():
        n, nq = len(pts), len(q)
        ridx, rdis = oracle.nn(pts, q)
        idx = np.zeros(nq, np.int32); dis = np.zeros(nq, np.float32)
        assert kd.kdr_nn(pts.ctypes.data_as(P), n, q.ctypes.data_as(P), nq, idx.ctypes.data_as(P), dis.ctypes.data_as(P)) == 0
        assert (idx == ridx).all() and (dis == rdis).all()
        built = C.c_int(0)
        idx[:] = -7
        assert kd.kdr_nn_lazy(pts.ctypes.data_as(P), n, q.ctypes.data_as(P), nq, idx.ctypes.data_as(P),
                              dis.ctypes.data_as(P), C.byref(built)) == 0
        assert (idx == ridx).all() and (dis == rdis).all()
        assert built.value <= oracle.tree(pts)[0]        # the lazy tree never holds more nodes than the full one


@pytest.mark.parametrize("cap", [1 << 16, 300, (1 << 16) + 1, 301])   # odd: sequential partition
def test_cooperative_lazy_search_matches_reference(kd, oracle, cap):
    """The L2 kernel's cooperative split (staged gather, sequential mean + partition, parallel child boxes)."""
    for pts, q in clouds():
        n, nq = len(pts), len(q)
        ridx, rdis = oracle.nn(pts, q)
        idx = np.full(nq, -7, np.int32); dis = np.zeros(nq, np.float32)
        built = C.c_int(0)
        assert kd.kdr_nn_lazy_coop(pts.ctypes.data_as(P), n, q.ctypes.data_as(P), nq, idx.ctypes.data_as(P),
                                   dis.ctypes.data_as(P), cap, C.byref(built)) == 0
        assert (idx == ridx).all() and (dis == rdis).all()


@pytest.mark.parametrize("cap", [300, 64, 301])   # odd: sequential partition
def test_cooperative_split_through_the_second_staging_pair(kd, oracle, cap):
    """Ranges beyond the (shared-memory) staging capacity are staged in the larger second pair (global memory on the device)
    instead of being split by one thread: same tree, same answers."""
    for pts, q in clouds():
        n, nq = len(pts), len(q)
        ridx, rdis = oracle.nn(pts, q)
        idx = np.full(nq, -7, np.int32); dis = np.zeros(nq, np.float32)
        built = C.c_int(0)
        assert kd.kdr_nn_lazy_coop_g(pts.ctypes.data_as(P), n, q.ctypes.data_as(P), nq, idx.ctypes.data_as(P),
                                     dis.ctypes.data_as(P), cap, n + 16, C.byref(built)) == 0
        assert (idx == ridx).all() and (dis == rdis).all()


def test_rank_table_partition_equals_the_sequential_partition(kd):
    """part_* (kdreplay.h): the data-parallel form of select_on_coordinate_value writes the very permutation the
    reference's swap loop (kdtree.cpp:294-317) produces -- every n up to 40 with random patterns, all 2^n patterns up
    to n = 10, long arrays, all-kept / all-moved / single-element edge cases, NaN, 1..1024 emulated threads."""
    rng = np.random.default_rng(5)
    U16 = C.POINTER(C.c_uint16)
    def check(v, alpha, nt):
        v = np.ascontiguousarray(v, np.float32)
        n = len(v)
        a = np.zeros(max(n, 1), np.uint16); b = np.zeros(max(n, 1), np.uint16)
        ms = C.c_int(-9)
        mp = kd.kdr_partition_both(v.ctypes.data_as(C.POINTER(C.c_float)), n, C.c_float(alpha), nt,
                                   a.ctypes.data_as(U16), b.ctypes.data_as(U16), C.byref(ms))
        assert mp == ms.value, (n, nt, v.tolist())
        assert np.array_equal(a[:n], b[:n]), (n, nt, v.tolist(), a[:n].tolist(), b[:n].tolist())
    for n in range(1, 11):
        for bits in range(1 << n):
            v = np.array([(bits >> i) & 1 for i in range(n)], np.float32)
            check(v, 0.5, 1); check(v, 0.5, 3)
    for n in range(1, 41):
        for _ in range(30):
            check(rng.integers(0, 2, n), 0.5, int(rng.integers(1, 9)))
    for n in (1000, 4097, 32768, 65535):
        for frac in (0.5, 0.02, 0.98):
            v = (rng.random(n) < frac).astype(np.float32)
            for nt in (1, 32, 1024):
                check(v, 0.5, nt)
        check(np.zeros(n), 0.5, 1024); check(np.ones(n), 0.5, 1024)
        w = rng.normal(size=n).astype(np.float32)
        check(w, float(w.mean()), 1024)
    w = rng.normal(size=500).astype(np.float32); w[::7] = np.nan
    check(w, 0.0, 64)
