#!/usr/bin/env python
"""Generates tests/golden/ccicp_golden.json from the UNMODIFIED reference sources compiled as
oracle/_ref/libccicp_ref.so (oracle/Makefile).  The reference ships no golden vectors or tests of its own
(SURVEY.md section 4), so these fixtures -- outputs of the reference itself run in the build container on
seeded synthetic scans -- are what pins the oracle and the CUDA path.  Re-run only where /root/reference
exists:   python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind          # noqa: E402
from slam_b200 import scangen       # noqa: E402

CASES = [  # (name, seed, config, class_rule, max_iter, min_delta, indist)
    ("c1_seed1", 1, 1, 0, 30, -2.0, 5.0),
    ("c1_seed1_all_nga", 1, 1, 1, 30, -2.0, 5.0),
    ("c1_seed1_all_ga", 1, 1, 2, 30, -2.0, 5.0),
    ("c2_seed1000", 1000, 1, 0, 30, -2.0, 5.0),
    ("c2_seed1001_defaults", 1001, 1, 0, 20, 1e-6, 5.0),
    ("c3_seed1000000", 10**6, 3, 0, 30, -2.0, 5.0),
    ("c1_seed9_tight_gate", 9, 1, 0, 12, -2.0, 0.04),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def hexf(a):
    return [float(x).hex() for x in np.asarray(a, np.float64).ravel()]


def lattice_case():
    rng = np.random.default_rng(5)
    model = np.round(rng.uniform(-12, 12, (900, 2)))
    scene = np.round(rng.uniform(-12, 12, (700, 2)) * 2) / 2
    return (model[:300], model[300:], scene[:250], scene[250:])


def main():
    assert refbind.available("reference"), "needs oracle/_ref (run `make -C oracle ref` where /root/reference exists)"
    out = {"generator": "tests/golden/make_golden.py", "oracle": "oracle/_ref (unmodified reference sources)", "cases": {}}
    for name, seed, cfg, rule, it, md, ind in CASES:
        b = scangen.generate(seed, 1, params=scangen.default_params(cfg, class_rule=rule))
        tr = refbind.trace(*b.pair(0), b.init_Rt()[0], max_iter=it, min_delta=md, indist=ind, want_info=True,
                           kind="reference")
        out["cases"][name] = {
            "seed": seed, "config": cfg, "class_rule": rule, "max_iter": it, "min_delta": md, "indist": ind,
            "input_sha256": sha(np.concatenate([b.model_xy.ravel(), b.scene_xy.ravel()])),
            "counts": [int(len(a)) for a in b.pair(0)],
            "iters": int(tr["iters"]), "ncorr": [int(x) for x in tr["ncorr"]],
            "Rt_final": hexf(tr["Rt"]), "Rt_iter1": hexf(tr["Rt_trace"][1]), "delta": hexf(tr["delta"]),
            "corr_sha256": sha(tr["corr"]), "corr_last_sha256": sha(tr["corr"][-1]), "dis_sha256": sha(tr["dis"]),
            "info": hexf(tr["info"]), "ties": int(tr["tie"].sum()),
        }
    pair = lattice_case()
    I6 = np.array([1.0, 0, 0, 1, 0, 0])
    tr = refbind.trace(*pair, I6, max_iter=10, min_delta=-2.0, kind="reference")
    out["cases"]["lattice_ties"] = {"iters": int(tr["iters"]), "ncorr": [int(x) for x in tr["ncorr"]],
                                    "ties_per_iter": [int(x) for x in tr["tie"].sum(1)],
                                    "corr_sha256": sha(tr["corr"]), "Rt_final": hexf(tr["Rt"])}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ccicp_golden.json")
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
