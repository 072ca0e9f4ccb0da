"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares; struct layouts match
the Python mirror; compute entry points fail loudly (never fall back) when there is no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from slam_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "ccicp2d_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ccicp_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = capi.load()
    names = declared_functions()
    assert len(names) >= 12 and set(names) == set(capi.SYMBOLS)
    for n in names:
        assert hasattr(lib, n), n
    assert lib.ccicp_abi_version() == 5


def test_struct_layouts():
    assert capi.RESULT_DTYPE.itemsize == 160
    assert capi.RESULT_DTYPE.fields["info"][1] == 48 and capi.RESULT_DTYPE.fields["num_corr"][1] == 136
    assert C.sizeof(capi.Params) == 88
    p = capi.default_params()
    assert (p.max_iter, p.min_delta, p.indist, p.flags) == (20, 1e-6, 5.0, 0)      # icp.cpp:27, icpTools.cpp:188


def test_error_strings_and_argument_checks():
    lib = capi.load()
    assert b"no CPU path" in lib.ccicp_strerror(capi.E_NO_DEVICE)
    assert lib.ccicp_create(None, 0) == capi.E_INVALID
    assert lib.ccicp_register_batch(None, 1, 1, None, None, None, None, None, None, None, None) == capi.E_INVALID
    assert lib.ccicp_destroy(None) == capi.OK


def test_no_device_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.CcicpError) as e:
        capi.Context(0)
    assert e.value.code == capi.E_NO_DEVICE


def test_facade_library_exports_reference_class_surface():
    """libccicp2d.so carries the reference's C++ symbols (Icp::Icp, Icp::fit, getEdgeWeight, ...)."""
    import subprocess
    lib = os.path.join(ROOT, "slam_b200", "lib", "libccicp2d.so")
    assert os.path.exists(lib)
    syms = subprocess.run(["nm", "-DC", lib], stdout=subprocess.PIPE, text=True).stdout
    for want in ["Icp::Icp(double*, double*, int, int, int)", "Icp::fit(double*, double*, int, int, Matrix&, Matrix&, double, double)",
                 "Icp::getEdgeWeight(double*)", "Icp::getNumberCorrespondences()"]:
        assert want in syms, want


def test_knn_candidates_follow_graph_slam_rules():
    """graphSlamGetKNN (graphSlamTools.cpp:72-106): last two nodes never considered, K clipped, self skipped."""
    xy = np.array([[0, 0], [1, 0], [2, 0], [3, 0], [10, 0], [0.5, 0]], float)
    q = capi.knn_candidates(xy, xy[5], 5, K=3)
    assert q.tolist() == [0, 1, 2]                      # nodes 4 and 5 are outside the i < n-2 loop; ties 0/1 -> lower index
    assert capi.knn_candidates(xy, xy[1], 1, K=3).tolist() == [0, 2]          # itself is skipped but still uses a slot
    assert capi.knn_candidates(xy[:3], xy[2], 2, K=3).tolist() == [0]        # K clipped to n-1 = 2, only node 0 in range
    assert capi.knn_candidates(xy[:2], xy[1], 1, K=3).tolist() == []
