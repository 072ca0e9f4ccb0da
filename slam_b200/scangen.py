"""ctypes binding of the seeded synthetic scan-pair generator (csrc/scangen.c):
workload generator for tests and bench.py."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libscangen.so")


class ScangenParams(C.Structure):
    _fields_ = [("n_beams", C.c_int32), ("max_boxes", C.c_int32), ("class_rule", C.c_int32), ("reserved", C.c_int32),
                ("fov_start_deg", C.c_double), ("fov_step_deg", C.c_double), ("noise_sigma", C.c_double),
                ("max_range", C.c_double), ("room_rmin", C.c_double), ("room_rmax", C.c_double),
                ("off_xy", C.c_double), ("off_th", C.c_double), ("init_sigma_xy", C.c_double),
                ("init_sigma_th", C.c_double), ("jump_thresh", C.c_double)]


_lib = None


def load():
    global _lib
    if _lib is None:
        lib = C.CDLL(LIB_PATH)
        lib.scangen_default.argtypes = [C.POINTER(ScangenParams), C.c_int32]
        lib.scangen_default.restype = None
        lib.scangen_batch.argtypes = [C.POINTER(ScangenParams), C.c_uint64, C.c_int32] + [C.c_void_p] * 5
        lib.scangen_batch.restype = C.c_int32
        lib.scangen_hypotheses.argtypes = [C.c_uint64, C.c_int32, C.c_void_p]
        lib.scangen_hypotheses.restype = None
        lib.scangen_count_float_duplicates.argtypes = [C.c_void_p, C.c_int32]
        lib.scangen_count_float_duplicates.restype = C.c_int32
        _lib = lib
    return _lib


def default_params(config=1, **kw):
    p = ScangenParams()
    load().scangen_default(C.byref(p), config)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def pose_to_Rt(pose):
    """[n,3] (x, y, yaw) -> [n,6] {r00,r01,r10,r11,t0,t1}; R = rot(yaw) as icpTools.cpp:168-176."""
    pose = np.asarray(pose, np.float64).reshape(-1, 3)
    c, s = np.cos(pose[:, 2]), np.sin(pose[:, 2])
    return np.ascontiguousarray(np.stack([c, -s, s, c, pose[:, 0], pose[:, 1]], 1))


def Rt_to_pose(Rt):
    """[n,6] -> [n,3] (x, y, yaw = atan2(R10, R00)) as icpTools.cpp:197."""
    Rt = np.asarray(Rt, np.float64).reshape(-1, 6)
    return np.stack([Rt[:, 4], Rt[:, 5], np.arctan2(Rt[:, 2], Rt[:, 0])], 1)


class Batch:
    """CSR-packed scan pairs: model_xy/scene_xy float64 [N,2], offsets int32 [2P+1], truth/init [P,3]."""

    def __init__(self, model_xy, model_off, scene_xy, scene_off, truth, init, seeds):
        self.model_xy, self.model_off, self.scene_xy, self.scene_off = model_xy, model_off, scene_xy, scene_off
        self.truth, self.init, self.seeds = truth, init, seeds

    @property
    def n_pairs(self):
        return (len(self.model_off) - 1) // 2

    def pair(self, p):
        mo, so = self.model_off, self.scene_off
        return (self.model_xy[mo[2 * p]:mo[2 * p + 1]], self.model_xy[mo[2 * p + 1]:mo[2 * p + 2]],
                self.scene_xy[so[2 * p]:so[2 * p + 1]], self.scene_xy[so[2 * p + 1]:so[2 * p + 2]])

    @property
    def max_model(self):
        return int((self.model_off[2::2] - self.model_off[:-2:2]).max()) if self.n_pairs else 0

    @property
    def max_scene(self):
        return int((self.scene_off[2::2] - self.scene_off[:-2:2]).max()) if self.n_pairs else 0

    def init_Rt(self):
        return pose_to_Rt(self.init)


def _pack(padded, n_a, n_b, out=None):
    """[P, B, 2] padded blocks with n_a+n_b valid leading rows -> CSR ([N,2], off[2P+1])."""
    P, B = padded.shape[0], padded.shape[1]
    tot = (n_a + n_b).astype(np.int64)
    off = np.zeros(2 * P + 1, np.int64)
    off[1::2] = n_a
    off[2::2] = n_b
    off = np.cumsum(off)
    off[0] = 0
    mask = np.arange(B)[None, :] < tot[:, None]
    xy = padded[mask]
    if out is not None:
        out[:len(xy)] = xy
        xy = out[:len(xy)]
    return np.ascontiguousarray(xy), off.astype(np.int32)


def generate(seed0, n_pairs, config=1, params=None, pinned=None):
    """Pairs with seeds seed0..seed0+n_pairs-1.  `pinned(shape, dtype)` optionally allocates the xy arrays."""
    p = params or default_params(config)
    lib = load()
    B = p.n_beams
    m = np.zeros((n_pairs, B, 2), np.float64)
    s = np.zeros((n_pairs, B, 2), np.float64)
    counts = np.zeros((n_pairs, 4), np.int32)
    truth = np.zeros((n_pairs, 3), np.float64)
    init = np.zeros((n_pairs, 3), np.float64)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = lib.scangen_batch(C.byref(p), seed0, n_pairs, vp(m), vp(s), vp(counts), vp(truth), vp(init))
    if rc != 0:
        raise ValueError("scangen_batch failed")
    mo = so = None
    if pinned is not None:
        mo = pinned((int(counts[:, :2].sum()), 2), np.float64)
        so = pinned((int(counts[:, 2:].sum()), 2), np.float64)
    mxy, moff = _pack(m, counts[:, 0], counts[:, 1], mo)
    sxy, soff = _pack(s, counts[:, 2], counts[:, 3], so)
    return Batch(mxy, moff, sxy, soff, truth, init, np.arange(seed0, seed0 + n_pairs, dtype=np.uint64))


def hypotheses(pair_seed, n_seeds):
    out = np.zeros((n_seeds, 3), np.float64)
    load().scangen_hypotheses(int(pair_seed), n_seeds, out.ctypes.data_as(C.c_void_p))
    return out


def count_float_duplicates(xy):
    xy = np.ascontiguousarray(xy, np.float64)
    return int(load().scangen_count_float_duplicates(xy.ctypes.data_as(C.c_void_p), len(xy)))
