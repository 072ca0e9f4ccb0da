"""ctypes binding of include/ccicp2d_b200.h (libccicp2d_b200.so).

The product path has no CPU fallback: if the library is missing, or there is no
sm_100 device, these calls raise.  The host mirror of the reference's C++ interface for the
path is the facade library (include/ccicp2d/*.h, slam_b200/csrc/facade/icp_facade.cpp); this
module is test / bench orchestration only.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CCICP_LIB") or os.path.join(_HERE, "lib", "libccicp2d_b200.so")

# error codes / status bits / flags (mirror the header)
OK, E_INVALID, E_NO_DEVICE, E_CUDA, E_NOMEM, E_TOO_LARGE, E_ARCH = range(7)
ST_TOO_FEW_MODEL, ST_TOO_FEW_SCENE, ST_NO_CORR, ST_REFLECTION, ST_DEGENERATE = 1, 2, 4, 8, 16
ST_TIE_REPLAY, ST_CONVERGED, ST_TREE_OVERFLOW, ST_INFO_SINGULAR, ST_HINT_TOO_SMALL = 32, 64, 128, 256, 512
ST_NO_SEED_BELOW, ST_EDGE_REJECTED, ST_EXACT_RETRY = 1024, 2048, 4096
STATIC_SCRATCH_BYTES, STATIC_MAX_POINTS, STATIC_MAX_PAIRS = 32 << 20, 2048, 65536   # ccicp_register_batch_device_static
F_INFO, F_INFO_FIXED_DY, F_BEST_OF_SEEDS, F_NO_TIE_REPLAY, F_EXACT_SUMS, F_FIRST_BELOW, F_TRIM = 1, 2, 4, 8, 16, 32, 64
F_DIAG_COUNTS, F_NO_EXACT_RETRY, F_CURV_GATE = 128, 256, 512


class Params(C.Structure):
    _fields_ = [("max_iter", C.c_int32), ("flags", C.c_uint32), ("min_delta", C.c_double),
                ("indist", C.c_double), ("min_corr_frac", C.c_double), ("score_threshold", C.c_double),
                ("reject_dist", C.c_double), ("reject_rot", C.c_double), ("trim_frac", C.c_double),
                ("curv_tol", C.c_double), ("feat_radius", C.c_double), ("feat_k", C.c_int32), ("reserved", C.c_int32)]


class LaunchInfo(C.Structure):
    _fields_ = [("grid", C.c_int32), ("block", C.c_int32), ("smem_bytes", C.c_int32),
                ("regs_per_thread", C.c_int32), ("blocks_per_sm", C.c_int32), ("kernel_variant", C.c_int32),
                ("n_launches", C.c_int32), ("sm_count", C.c_int32), ("last_kernel_ms", C.c_float),
                ("last_h2d_ms", C.c_float), ("last_d2h_ms", C.c_float)]


RESULT_DTYPE = np.dtype([("R", "<f8", (4,)), ("t", "<f8", (2,)), ("info", "<f8", (9,)), ("last_delta", "<f8"),
                         ("score", "<f8"), ("num_corr", "<i4"), ("iters", "<i4"), ("status", "<u4"),
                         ("best_seed", "<i4"), ("n_tie_replays", "<i4"), ("reserved", "<i4")])
assert RESULT_DTYPE.itemsize == 160

SYMBOLS = ["ccicp_abi_version", "ccicp_create", "ccicp_destroy", "ccicp_strerror", "ccicp_last_cuda_error",
           "ccicp_default_params", "ccicp_register_batch", "ccicp_register_batch_device", "ccicp_stream_sync",
           "ccicp_host_alloc", "ccicp_host_free", "ccicp_get_launch_info", "ccicp_match_batch", "ccicp_pose_to_rt",
           "ccicp_rt_to_pose", "ccicp_classify_points", "ccicp_classify_points_device", "ccicp_register_batch_device_ex",
           "ccicp_grid_cells", "ccicp_edge_information_lum", "ccicp_knn_candidates", "ccicp_get_diag_counts",
           "ccicp_multi_create", "ccicp_multi_destroy", "ccicp_multi_device_count", "ccicp_multi_ctx", "ccicp_multi_last_error",
           "ccicp_shard_range", "ccicp_register_batch_multi", "ccicp_register_batch_multi_device", "ccicp_multi_sync",
           "ccicp_multi_gather_kind", "ccicp_scans_upload", "ccicp_register_scan_pairs", "ccicp_scan_features",
           "ccicp_register_batch_device_static"]

_lib = None


class CcicpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("ccicp error %d: %s" % (code, msg))
        self.code = code


def load():
    """Loads the CUDA library; raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: run `python -m slam_b200.build` (there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, dbl = C.c_void_p, C.c_int32, C.c_double
    lib.ccicp_abi_version.restype = C.c_int
    lib.ccicp_create.argtypes = [C.POINTER(vp), C.c_int]
    lib.ccicp_destroy.argtypes = [vp]
    lib.ccicp_strerror.argtypes = [C.c_int]
    lib.ccicp_strerror.restype = C.c_char_p
    lib.ccicp_last_cuda_error.argtypes = [vp]
    lib.ccicp_last_cuda_error.restype = C.c_char_p
    lib.ccicp_default_params.argtypes = [C.POINTER(Params)]
    lib.ccicp_default_params.restype = None
    lib.ccicp_register_batch.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp]
    lib.ccicp_register_batch_device.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp, i32, i32, vp]
    lib.ccicp_register_batch_device_ex.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp, i32, i32, i32, vp]
    lib.ccicp_register_batch_device_static.argtypes = [vp, i32, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp, i32, i32, vp]
    lib.ccicp_edge_information_lum.argtypes = [vp, i32, vp, vp, vp, vp, dbl, vp, vp, vp]
    lib.ccicp_knn_candidates.argtypes = [vp, i32, vp, i32, i32, vp]
    lib.ccicp_knn_candidates.restype = i32
    lib.ccicp_grid_cells.argtypes = [i32]
    lib.ccicp_grid_cells.restype = i32
    lib.ccicp_stream_sync.argtypes = [vp, vp]
    lib.ccicp_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    lib.ccicp_host_free.argtypes = [vp]
    lib.ccicp_get_launch_info.argtypes = [vp, C.POINTER(LaunchInfo)]
    lib.ccicp_get_diag_counts.argtypes = [vp, vp]
    lib.ccicp_scan_features.argtypes = [vp, i32, vp, vp, i32, dbl, vp, vp]
    lib.ccicp_scans_upload.argtypes = [vp, i32, vp, vp]
    lib.ccicp_register_scan_pairs.argtypes = [vp, i32, i32, vp, vp, vp, C.POINTER(Params), vp]
    lib.ccicp_match_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp]
    lib.ccicp_classify_points.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    lib.ccicp_classify_points_device.argtypes = [vp, i32, vp, vp, C.c_int64, vp, vp, vp, vp, vp]
    lib.ccicp_pose_to_rt.argtypes = [vp, vp]
    lib.ccicp_pose_to_rt.restype = None
    lib.ccicp_rt_to_pose.argtypes = [vp, vp, vp]
    lib.ccicp_rt_to_pose.restype = None
    _lib = lib
    return lib


def default_params(**kw):
    p = Params()
    load().ccicp_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def fixed_iter_params(iters=30, flags=0, indist=5.0, trim_frac=1.0):
    """BASELINE.json's setting: setMaxIterations(30) + setMinDeltaParam(-2)."""
    return default_params(max_iter=iters, min_delta=-2.0, flags=flags, indist=indist, trim_frac=trim_frac)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """One ccicp_ctx: a device, a stream, grow-only device buffers."""

    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.ccicp_create(C.byref(h), device)
        if rc != OK:
            raise CcicpError(rc, self.lib.ccicp_strerror(rc).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.ccicp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != OK:
            msg = self.lib.ccicp_strerror(rc).decode()
            extra = self.lib.ccicp_last_cuda_error(self.h).decode()
            raise CcicpError(rc, msg + (" [" + extra + "]" if extra else ""))

    def register_batch(self, model_xy, model_off, scene_xy, scene_off, init_Rt, params, n_seeds=1, want_corr=False,
                       out=None, corr_out=None):
        """Host-buffer call (ccicp_register_batch).  Arrays: float64 xy, int32 CSR offsets, float64 [n,6]."""
        model_off = np.ascontiguousarray(model_off, np.int32)
        scene_off = np.ascontiguousarray(scene_off, np.int32)
        n_pairs = (len(model_off) - 1) // 2
        model_xy = np.ascontiguousarray(model_xy, np.float64)
        scene_xy = np.ascontiguousarray(scene_xy, np.float64)
        init_Rt = np.ascontiguousarray(init_Rt, np.float64)
        assert init_Rt.size == n_pairs * n_seeds * 6
        best_of = bool(params.flags & F_BEST_OF_SEEDS) and n_seeds > 1
        n_out = n_pairs if best_of else n_pairs * n_seeds
        if out is None:
            out = np.zeros(n_out, RESULT_DTYPE)
        corr = None
        if want_corr:
            corr = corr_out if corr_out is not None else np.full((n_seeds, int(scene_off[-1])), -2, np.int32)
        rc = self.lib.ccicp_register_batch(self.h, n_pairs, n_seeds, _ptr(model_xy), _ptr(model_off), _ptr(scene_xy),
                                           _ptr(scene_off), _ptr(init_Rt), C.byref(params), _ptr(out), _ptr(corr))
        self._check(rc)
        return (out, corr) if want_corr else out

    def match_batch(self, model_xy, model_off, scene_xy, scene_off, init_pose, params, want_records=False):
        """Batched CCICP::doICPMatch (ccicp_match_batch).  init_pose: float64 [n,7] = x,y,z,qx,qy,qz,qw."""
        model_off = np.ascontiguousarray(model_off, np.int32)
        scene_off = np.ascontiguousarray(scene_off, np.int32)
        n_pairs = (len(model_off) - 1) // 2
        model_xy = np.ascontiguousarray(model_xy, np.float64)
        scene_xy = np.ascontiguousarray(scene_xy, np.float64)
        init_pose = np.ascontiguousarray(init_pose, np.float64).reshape(n_pairs, 7)
        out = np.zeros((n_pairs, 7), np.float64)
        rec = np.zeros(n_pairs, RESULT_DTYPE) if want_records else None
        rc = self.lib.ccicp_match_batch(self.h, n_pairs, _ptr(model_xy), _ptr(model_off), _ptr(scene_xy), _ptr(scene_off),
                                        _ptr(init_pose), C.byref(params), _ptr(out), _ptr(rec))
        self._check(rc)
        return (out, rec) if want_records else out

    def edge_information_lum(self, source_clouds, target_clouds, max_corr_distance=0.75):
        """Batched computeEdgeInformationLUM.  Lists of float32 [n,3] clouds (sources already transformed).
        Returns (info [E,6,6], num_corr [E], status [E])."""
        E = len(source_clouds)
        cat = lambda cl: np.ascontiguousarray(np.concatenate([np.asarray(c, np.float32).reshape(-1, 3) for c in cl], 0)
                                              if cl else np.zeros((0, 3), np.float32))
        off = lambda cl: np.cumsum([0] + [len(c) for c in cl]).astype(np.int32)
        s, t, so, to = cat(source_clouds), cat(target_clouds), off(source_clouds), off(target_clouds)
        info = np.zeros((max(E, 1), 6, 6), np.float64); num = np.zeros(max(E, 1), np.int32); st = np.zeros(max(E, 1), np.uint32)
        self._check(self.lib.ccicp_edge_information_lum(self.h, E, _ptr(s), _ptr(so), _ptr(t), _ptr(to), float(max_corr_distance),
                                                        _ptr(info), _ptr(num), _ptr(st)))
        return info[:E], num[:E], st[:E]

    def classify_points(self, xyz, cloud_off):
        """Batched CCICP::classifyPoints.  xyz float32 [N,3], cloud_off int32 [n_clouds+1].
        Returns a list of (xyz[k,3], ground_adj[k], src_index[k]) per cloud."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        cloud_off = np.ascontiguousarray(cloud_off, np.int32)
        nc, n = len(cloud_off) - 1, len(xyz)
        oxyz = np.zeros((max(n, 1), 3), np.float32); oga = np.zeros(max(n, 1), np.uint8)
        osrc = np.zeros(max(n, 1), np.int32); ocnt = np.zeros(max(nc, 1), np.int32)
        self._check(self.lib.ccicp_classify_points(self.h, nc, _ptr(xyz), _ptr(cloud_off), _ptr(oxyz), _ptr(oga),
                                                   _ptr(osrc), _ptr(ocnt)))
        out = []
        for c in range(nc):
            a, k = int(cloud_off[c]), int(ocnt[c])
            out.append((oxyz[a:a + k].copy(), oga[a:a + k].copy(), osrc[a:a + k].copy()))
        return out

    def classify_points_device(self, n_clouds, d_xyz, d_off, n_points, d_out_xyz, d_out_ga, d_out_src, d_out_cnt, stream=None):
        self._check(self.lib.ccicp_classify_points_device(self.h, n_clouds, d_xyz, d_off, n_points, d_out_xyz, d_out_ga,
                                                          d_out_src or None, d_out_cnt, stream or None))

    def register_batch_device(self, n_pairs, n_seeds, d_model_xy, d_model_off, d_scene_xy, d_scene_off, d_init_Rt,
                              params, d_out, d_corr, max_model, max_scene, stream=None, max_cells_hint=0):
        """Device-pointer call (ints from tensor.data_ptr()); asynchronous on `stream` (cudaStream_t as int)."""
        rc = self.lib.ccicp_register_batch_device_ex(self.h, n_pairs, n_seeds, d_model_xy, d_model_off, d_scene_xy,
                                                     d_scene_off, d_init_Rt, C.byref(params), d_out, d_corr or None,
                                                     max_model, max_scene, max_cells_hint, stream or None)
        self._check(rc)

    def register_batch_device_static(self, n_pairs, d_model_xy, d_model_off, d_scene_xy, d_scene_off, d_init_Rt, params,
                                     d_out, d_corr, max_model, max_scene, stream=None):
        """ccicp_register_batch_device_static: the allocation-free device-pointer call (scratch in a fixed device array
        of STATIC_SCRATCH_BYTES; n_pairs <= STATIC_MAX_PAIRS, scans <= STATIC_MAX_POINTS points)."""
        rc = self.lib.ccicp_register_batch_device_static(self.h, n_pairs, d_model_xy, d_model_off, d_scene_xy, d_scene_off,
                                                         d_init_Rt, C.byref(params), d_out, d_corr or None, max_model,
                                                         max_scene, stream or None)
        self._check(rc)

    def sync(self, stream=None):
        self._check(self.lib.ccicp_stream_sync(self.h, stream or None))

    def scan_features(self, xy, off, feat_k=5, feat_radius=0.5, want_normals=True):
        """ccicp_scan_features: per-point (normal [n,2], curvature [n]) of clouds in the registration calls' CSR layout."""
        off = np.ascontiguousarray(off, np.int32)
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
        n = len(xy)
        nrm = np.zeros((max(n, 1), 2), np.float32) if want_normals else None
        cur = np.zeros(max(n, 1), np.float32)
        self._check(self.lib.ccicp_scan_features(self.h, (len(off) - 1) // 2, _ptr(xy), _ptr(off), int(feat_k), float(feat_radius),
                                                 _ptr(nrm), _ptr(cur)))
        return (nrm[:n] if want_normals else None), cur[:n]

    def scans_upload(self, scan_xy, scan_off):
        """ccicp_scans_upload: the ctx's scan table (float64 xy, int32 offsets [2*n_scans+1])."""
        scan_off = np.ascontiguousarray(scan_off, np.int32)
        scan_xy = np.ascontiguousarray(scan_xy, np.float64)
        self._check(self.lib.ccicp_scans_upload(self.h, (len(scan_off) - 1) // 2, _ptr(scan_xy), _ptr(scan_off)))

    def register_scan_pairs(self, model_scan, scene_scan, init_Rt, params, n_seeds=1, out=None):
        """ccicp_register_scan_pairs: pair p registers scan scene_scan[p] against scan model_scan[p] of the table."""
        model_scan = np.ascontiguousarray(model_scan, np.int32)
        scene_scan = np.ascontiguousarray(scene_scan, np.int32)
        n_pairs = len(model_scan)
        init_Rt = np.ascontiguousarray(init_Rt, np.float64)
        assert init_Rt.size == n_pairs * n_seeds * 6 and len(scene_scan) == n_pairs
        best_of = bool(params.flags & F_BEST_OF_SEEDS) and n_seeds > 1
        if out is None:
            out = np.zeros(n_pairs if best_of else n_pairs * n_seeds, RESULT_DTYPE)
        self._check(self.lib.ccicp_register_scan_pairs(self.h, n_pairs, n_seeds, _ptr(model_scan), _ptr(scene_scan), _ptr(init_Rt),
                                                       C.byref(params), _ptr(out)))
        return out

    def diag_counts(self):
        """(distance evaluations, queries searched, accepted pairs) of the F_DIAG_COUNTS launches since the last call."""
        out = np.zeros(3, np.uint64)
        self._check(self.lib.ccicp_get_diag_counts(self.h, _ptr(out)))
        return int(out[0]), int(out[1]), int(out[2])

    def launch_info(self):
        li = LaunchInfo()
        self._check(self.lib.ccicp_get_launch_info(self.h, C.byref(li)))
        return li


def knn_candidates(nodes_xy, query_xy, query_idx, K=3):
    """graphSlamGetKNN: indices of the loop-closure candidates of a key frame (host logic)."""
    nodes_xy = np.ascontiguousarray(nodes_xy, np.float64).reshape(-1, 2)
    q = np.ascontiguousarray(query_xy, np.float64).reshape(2)
    out = np.zeros(max(K, 1), np.int32)
    n = load().ccicp_knn_candidates(_ptr(nodes_xy), len(nodes_xy), _ptr(q), int(query_idx), int(K), _ptr(out))
    return out[:n].copy()


def max_cells(model_off):
    """max over pairs of ccicp_grid_cells(nGA) + ccicp_grid_cells(nNGA): the hint of register_batch_device."""
    lib = load()
    off = np.asarray(model_off, np.int64)
    sizes = np.diff(off)
    cells = {int(n): lib.ccicp_grid_cells(int(n)) for n in np.unique(sizes)}
    per = np.array([cells[int(n)] for n in sizes], np.int64)
    return int((per[0::2] + per[1::2]).max()) if len(per) else 0


def pinned_array(shape, dtype):
    """numpy array backed by ccicp_host_alloc'd (page-locked) memory; keeps itself alive via .base."""
    lib = load()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    rc = lib.ccicp_host_alloc(C.byref(p), max(n, 1))
    if rc != OK:
        raise CcicpError(rc, lib.ccicp_strerror(rc).decode())
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[id(buf)] = (buf, p)
    return arr


_PINNED = {}


def pose_to_Rt(pose7):
    """ccicp_pose_to_rt on one pose (x,y,z,qx,qy,qz,qw) -> [6]."""
    p = np.ascontiguousarray(pose7, np.float64).reshape(7)
    out = np.zeros(6, np.float64)
    load().ccicp_pose_to_rt(_ptr(p), _ptr(out))
    return out


def Rt_to_pose(Rt6, init_pose7):
    r = np.ascontiguousarray(Rt6, np.float64).reshape(6)
    p = np.ascontiguousarray(init_pose7, np.float64).reshape(7)
    out = np.zeros(7, np.float64)
    load().ccicp_rt_to_pose(_ptr(r), _ptr(p), _ptr(out))
    return out
