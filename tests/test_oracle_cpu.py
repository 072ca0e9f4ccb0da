"""CPU tests of the checkers themselves (no GPU): the oracle is pinned against the committed golden vectors
(outputs of the unmodified reference, tests/golden/make_golden.py) and, where both are present, the C
restatement (oracle/ccicp_port.c) against the compiled reference bit for bit."""
import hashlib
import importlib.util
import json
import os

import numpy as np
import pytest

from slam_b200 import scangen

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "ccicp_golden.json")))
I6 = np.array([1.0, 0, 0, 1, 0, 0])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def unhex(l):
    return np.array([float.fromhex(x) for x in l])


def kinds(oracle):
    return [k for k in ("reference", "port") if oracle.available(k)]


@pytest.mark.parametrize("name", [k for k in GOLD["cases"] if k != "lattice_ties"])
def test_oracle_reproduces_golden(oracle, name):
    g = GOLD["cases"][name]
    b = scangen.generate(g["seed"], 1, params=scangen.default_params(g["config"], class_rule=g["class_rule"]))
    assert sha(np.concatenate([b.model_xy.ravel(), b.scene_xy.ravel()])) == g["input_sha256"], "scan generator drifted"
    assert [len(a) for a in b.pair(0)] == g["counts"]
    for kind in kinds(oracle):
        tr = oracle.trace(*b.pair(0), b.init_Rt()[0], max_iter=g["max_iter"], min_delta=g["min_delta"],
                          indist=g["indist"], want_info=True, kind=kind)
        assert tr["iters"] == g["iters"] and list(tr["ncorr"]) == g["ncorr"], kind
        assert sha(tr["corr"]) == g["corr_sha256"] and sha(tr["dis"]) == g["dis_sha256"], kind
        assert (tr["Rt"] == unhex(g["Rt_final"])).all(), kind           # bit for bit
        assert (tr["delta"] == unhex(g["delta"])).all(), kind
        assert (tr["info"] == unhex(g["info"])).all(), kind
        assert int(tr["tie"].sum()) == g["ties"]
        # the public-API entry point agrees with the stepped trace
        r = oracle.fit(*b.pair(0), b.init_Rt()[0], max_iter=g["max_iter"], min_delta=g["min_delta"],
                       indist=g["indist"], want_info=True, kind=kind)
        assert (r["Rt"] == tr["Rt"]).all() and r["num_corr"] == g["ncorr"][-1]


def test_oracle_tie_rule_golden(oracle):
    """Lattice clouds: most queries have 2-4 equidistant model points; the answer follows the kd-tree traversal."""
    g = GOLD["cases"]["lattice_ties"]
    rng = np.random.default_rng(5)
    model = np.round(rng.uniform(-12, 12, (900, 2)))
    scene = np.round(rng.uniform(-12, 12, (700, 2)) * 2) / 2
    pair = (model[:300], model[300:], scene[:250], scene[250:])
    for kind in kinds(oracle):
        tr = oracle.trace(*pair, I6, max_iter=10, min_delta=-2.0, kind=kind)
        assert list(tr["tie"].sum(1)) == g["ties_per_iter"] and sha(tr["corr"]) == g["corr_sha256"], kind
        assert (tr["Rt"] == unhex(g["Rt_final"])).all()


def test_port_equals_compiled_reference(oracle):
    if not (oracle.available("reference") and oracle.available("port")):
        pytest.skip("needs both oracle/_ref and oracle/libccicp_port.so")
    rng = np.random.default_rng(0)
    for seed, cfg in [(3, 1), (77, 3), (12345, 1)]:
        b = scangen.generate(seed, 1, config=cfg)
        a = oracle.trace(*b.pair(0), b.init_Rt()[0], want_info=True, kind="reference")
        c = oracle.trace(*b.pair(0), b.init_Rt()[0], want_info=True, kind="port")
        for k in ("Rt_trace", "corr", "dis", "delta", "tie", "info"):
            assert (a[k] == c[k]).all(), (seed, k)
    pts = np.round(rng.uniform(-40, 40, (1500, 2)))
    q = (np.round(rng.uniform(-42, 42, (3000, 2)) * 2) / 2).astype(np.float32)
    ia, da = oracle.nn(pts, q, kind="reference")
    ib, db = oracle.nn(pts, q, kind="port")
    assert (ia == ib).all() and (da == db).all()
    ta, tb = oracle.tree(pts, kind="reference"), oracle.tree(pts, kind="port")
    assert ta[0] == tb[0] and (ta[1] == tb[1]).all() and (ta[2].view(np.uint32) == tb[2].view(np.uint32)).all()


def test_oracle_edge_cases(oracle):
    rng = np.random.default_rng(1)
    big = rng.uniform(-5, 5, (50, 2))
    E = np.zeros((0, 2))
    for kind in kinds(oracle):
        Rt0 = np.array([0.8, -0.6, 0.6, 0.8, 1.5, -2.5])
        r = oracle.fit(big[:2], big[2:4], big[:20], big[20:], Rt0, kind=kind)
        assert r["status"] == oracle.TOO_FEW_MODEL and (r["Rt"] == Rt0).all()
        r = oracle.fit(big[:20], big[20:], big[:1], big[1:4], Rt0, kind=kind)
        assert r["status"] == oracle.TOO_FEW_SCENE and (r["Rt"] == Rt0).all()
        far = oracle.trace(E, big, E, big + 100.0, I6, max_iter=7, min_delta=1e-6, kind=kind)
        assert far["status"] == oracle.NO_CORR and far["iters"] == 1 and far["delta"][0] == -1.0
        far = oracle.trace(E, big, E, big + 100.0, I6, max_iter=7, min_delta=-2.0, kind=kind)
        assert far["iters"] == 7 and (far["Rt"] == I6).all()
        # a class whose model has <= 3 points is never searched
        tr = oracle.trace(big[:3], big[3:], big[:10], big[10:], I6, max_iter=2, min_delta=-2.0, kind=kind)
        assert (tr["corr"][:, :10] == -1).all() and (tr["corr"][:, 10:] >= 0).all()
