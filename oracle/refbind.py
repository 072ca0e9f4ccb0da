"""oracle/refbind.py -- TEST INFRASTRUCTURE ONLY.

ctypes binding of the CPU checkers:
  oracle/_ref/libccicp_ref.so  the UNMODIFIED reference sources + ref_probe.cpp ("reference")
  oracle/libccicp_port.so      the plain-C restatement ccicp_port.c ("port")
Both export the same ccref_* / ccport_* entry points.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product (slam_b200/) never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_PATH = os.path.join(_HERE, "_ref", "libccicp_ref.so")
PORT_PATH = os.path.join(_HERE, "libccicp_port.so")
CLASSIFY_REF_PATH = os.path.join(_HERE, "_ref", "libclassify_ref.so")

OK, TOO_FEW_MODEL, TOO_FEW_SCENE, NO_CORR = 0, 1, 2, 4
_libs = {}


def available(kind):
    return os.path.exists(REF_PATH if kind == "reference" else PORT_PATH)


def best_kind():
    return "reference" if available("reference") else "port"


def load(kind=None):
    kind = kind or best_kind()
    if kind not in _libs:
        path = REF_PATH if kind == "reference" else PORT_PATH
        lib = C.CDLL(path)
        pre = "ccref_" if kind == "reference" else "ccport_"
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        fit = getattr(lib, pre + "fit")
        fit.argtypes = [vp, i, vp, i, vp, i, vp, i, i, d, d, vp, vp, vp]
        fit.restype = i
        tr = getattr(lib, pre + "trace")
        tr.argtypes = [vp, i, vp, i, vp, i, vp, i, i, d, d, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        tr.restype = i
        tree = getattr(lib, pre + "tree")
        tree.argtypes = [vp, i, vp, vp, i]
        tree.restype = i
        nn = getattr(lib, pre + "nn")
        nn.argtypes = [vp, i, vp, i, vp, vp]
        nn.restype = None
        _libs[kind] = dict(fit=fit, trace=tr, tree=tree, nn=nn, lib=lib)
    return _libs[kind]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a):
    return np.ascontiguousarray(a, np.float64)


def fit(mGA, mNGA, sGA, sNGA, Rt0, max_iter=30, min_delta=-2.0, indist=5.0, want_info=False, kind=None):
    """One registration through the reference's public API.  Returns dict(Rt, num_corr, status, info)."""
    f = load(kind)["fit"]
    mGA, mNGA, sGA, sNGA = _c(mGA), _c(mNGA), _c(sGA), _c(sNGA)
    Rt = np.array(Rt0, np.float64).reshape(6).copy()
    nc = C.c_int(0)
    eW = np.zeros(9, np.float64)
    st = f(_p(mGA), len(mGA), _p(mNGA), len(mNGA), _p(sGA), len(sGA), _p(sNGA), len(sNGA), max_iter, min_delta,
           indist, _p(Rt), C.addressof(nc), _p(eW) if want_info else None)
    return dict(Rt=Rt, num_corr=nc.value, status=st, info=eW if want_info else None)


def trace(mGA, mNGA, sGA, sNGA, Rt0, max_iter=30, min_delta=-2.0, indist=5.0, want_info=False, kind=None, want_tie=True):
    """Lock-step trace.  Returns dict with Rt_trace[iters+1,6], corr[iters,ns], dis, tie, ncorr, delta, iters.
    want_tie=False skips the brute-force tie count (O(n_model) per query: minutes on 32768-point pairs)."""
    f = load(kind)["trace"]
    mGA, mNGA, sGA, sNGA = _c(mGA), _c(mNGA), _c(sGA), _c(sNGA)
    ns = len(sGA) + len(sNGA)
    Rt0 = np.array(Rt0, np.float64).reshape(6).copy()
    Rt_trace = np.zeros((max_iter + 1, 6), np.float64)
    corr = np.full((max_iter, max(ns, 1)), -3, np.int32)
    dis = np.zeros((max_iter, max(ns, 1)), np.float32)
    tie = np.zeros((max_iter, max(ns, 1)), np.uint8)
    ncorr = np.zeros(max_iter, np.int32)
    delta = np.zeros(max_iter, np.float64)
    iters = C.c_int(0)
    eW = np.zeros(9, np.float64)
    st = f(_p(mGA), len(mGA), _p(mNGA), len(mNGA), _p(sGA), len(sGA), _p(sNGA), len(sNGA), max_iter, min_delta,
           indist, _p(Rt0), _p(Rt_trace), _p(corr), _p(dis), _p(tie) if want_tie else None, _p(ncorr), _p(delta), C.addressof(iters),
           _p(eW) if want_info else None)
    k = iters.value
    return dict(status=st, iters=k, Rt_trace=Rt_trace[:k + 1], corr=corr[:k, :ns], dis=dis[:k, :ns],
                tie=tie[:k, :ns], ncorr=ncorr[:k], delta=delta[:k], info=eW if want_info else None,
                Rt=Rt_trace[k].copy())


def tree(pts, kind=None):
    f = load(kind)["tree"]
    pts = _c(pts)
    n = len(pts)
    ind = np.zeros(max(n, 1), np.int32)
    nodes = np.zeros(10 * (2 * n + 2), np.float32)
    nn = f(_p(pts), n, _p(ind), _p(nodes), 2 * n + 2)
    return nn, ind[:n], nodes[:10 * max(nn, 0)].reshape(-1, 10)


def nn(pts, q, kind=None):
    f = load(kind)["nn"]
    pts = _c(pts)
    q = np.ascontiguousarray(q, np.float32)
    idx = np.zeros(len(q), np.int32)
    dis = np.zeros(len(q), np.float32)
    f(_p(pts), len(pts), _p(q), len(q), _p(idx), _p(dis))
    return idx, dis


def classify_available():
    return os.path.exists(CLASSIFY_REF_PATH)


def classify_ref(xyz):
    """The reference's own CCICP::classifyPoints (icpTools.cpp:36-103, compiled unmodified: oracle/Makefile).
    xyz float32 [n,3], finite coordinates only.  Returns (xyz_out [k,3], ground_adj [k])."""
    if "classify" not in _libs:
        lib = C.CDLL(CLASSIFY_REF_PATH)
        lib.ccref_classify.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.ccref_classify.restype = C.c_int
        _libs["classify"] = lib
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    n = len(xyz)
    out = np.zeros((max(n, 1), 3), np.float32)
    ga = np.zeros(max(n, 1), np.uint8)
    k = _libs["classify"].ccref_classify(_p(xyz), n, _p(out), _p(ga))
    return out[:k].copy(), ga[:k].copy()
