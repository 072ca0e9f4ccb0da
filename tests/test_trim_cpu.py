"""oracle/trim_port.py (CPU statement of the CCICP_F_TRIM extra): with trim_frac = 1 it must reproduce the reference's
own fit; with trim_frac < 1 the kept set has the stated size."""
import math

import numpy as np

from slam_b200 import scangen
from oracle import trim_port
from helpers import assert_pose_close


def test_trim_frac_one_is_the_reference(oracle):
    for seed in (5, 6, 7):
        b = scangen.generate(seed, 1, config=3)
        want = oracle.fit(*b.pair(0), b.init_Rt()[0], max_iter=12, min_delta=-2.0)
        got = trim_port.fit_trimmed(*b.pair(0), b.init_Rt()[0], max_iter=12, trim_frac=1.0)
        assert got["num_corr"] == want["num_corr"]
        assert_pose_close(got["Rt"], want["Rt"], "seed %d" % seed)


def test_trimmed_set_size():
    b = scangen.generate(8, 1, config=3)
    full = trim_port.fit_trimmed(*b.pair(0), b.init_Rt()[0], max_iter=1, trim_frac=1.0)
    part = trim_port.fit_trimmed(*b.pair(0), b.init_Rt()[0], max_iter=1, trim_frac=0.7)
    k = math.ceil(0.7 * full["num_corr"])
    assert k <= part["num_corr"] <= k + 2                     # == k unless distances tie at the threshold
    assert ((part["corr"] >= 0) <= (full["corr"] >= 0)).all()  # a subset of the gated set
    assert (part["corr"][part["corr"] >= 0] == full["corr"][part["corr"] >= 0]).all()
