#!/usr/bin/env python
"""What a C++ caller gets from the multi-device entry points of the C ABI (ONE process, one host thread + ctx per GPU
inside the library): C2 weak scaling, 4096 pairs per device slot.
  host path    ccicp_register_batch_multi        pinned host buffers in, records gathered into the caller's out[]
  device path  ccicp_register_batch_multi_device shards resident in HBM, every device ends with all records (NCCL / peer)
Timed with the host clock around K calls (+ ccicp_multi_sync): the calls span several devices, so there is no single
stream to put CUDA events on; the per-device kernel times come from each slot's ctx.
usage: python profiles/multi_bench.py 0,1,...   -> one JSON line"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from slam_b200 import capi, scangen, sharding  # noqa: E402
import bench  # noqa: E402

devs = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0").split(",")]
G, P, steps, warm = len(devs), 4096, 10, 3
lib = capi.load()
vp = C.c_void_p
lib.ccicp_multi_create.argtypes = [vp, C.c_int, C.POINTER(vp)]
lib.ccicp_multi_destroy.argtypes = [vp]
lib.ccicp_multi_sync.argtypes = [vp]
lib.ccicp_multi_gather_kind.argtypes = [vp]
lib.ccicp_multi_ctx.argtypes = [vp, C.c_int]
lib.ccicp_multi_ctx.restype = vp
lib.ccicp_register_batch_multi.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, C.POINTER(capi.Params), vp, vp]
lib.ccicp_register_batch_multi_device.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, C.POINTER(capi.Params), vp, vp, vp, vp]
m = vp()
assert lib.ccicp_multi_create((C.c_int * G)(*devs), G, C.byref(m)) == 0
n = G * P
b = bench.generate_chunked(bench.SEED0, n, 1, capi.pinned_array)
init = capi.pinned_array((n, 6), np.float64); init[:] = b.init_Rt()
out = capi.pinned_array((n,), capi.RESULT_DTYPE)
prm = capi.fixed_iter_params(bench.ITERS, indist=bench.INDIST)
ptr = lambda a: a.ctypes.data_as(vp)


def host_call():
    rc = lib.ccicp_register_batch_multi(m, n, 1, ptr(b.model_xy), ptr(b.model_off), ptr(b.scene_xy), ptr(b.scene_off), ptr(init),
                                        C.byref(prm), ptr(out), None)
    assert rc == 0, rc


for _ in range(warm):
    host_call()
t0 = time.perf_counter()
for _ in range(steps):
    host_call()
host_ms = (time.perf_counter() - t0) * 1e3 / steps
ref = out.copy()

# device-resident shards
keep, arrs = [], {k: [] for k in ("mx", "mo", "sx", "so", "in", "out")}
mm, ms, mc = [], [], []
for g, d in enumerate(devs):
    lo, hi = sharding.shard_range(n, g, G)
    dev = torch.device("cuda", d)
    mo = (b.model_off[2 * lo:2 * hi + 1] - b.model_off[2 * lo]).astype(np.int32)
    so = (b.scene_off[2 * lo:2 * hi + 1] - b.scene_off[2 * lo]).astype(np.int32)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
    ts = [t(b.model_xy[b.model_off[2 * lo]:b.model_off[2 * hi]], np.float64), t(mo, np.int32),
          t(b.scene_xy[b.scene_off[2 * lo]:b.scene_off[2 * hi]], np.float64), t(so, np.int32), t(init[lo:hi], np.float64),
          torch.zeros(n * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)]
    keep.append(ts)
    for k, x in zip(("mx", "mo", "sx", "so", "in", "out"), ts):
        arrs[k].append(x.data_ptr())
    mm.append(int((mo[2::2] - mo[:-2:2]).max())); ms.append(int((so[2::2] - so[:-2:2]).max())); mc.append(capi.max_cells(mo))
pa = {k: (vp * G)(*v) for k, v in arrs.items()}
i32 = lambda v: (C.c_int32 * G)(*v)
a_mm, a_ms, a_mc = i32(mm), i32(ms), i32(mc)


def dev_call():
    rc = lib.ccicp_register_batch_multi_device(m, n, 1, pa["mx"], pa["mo"], pa["sx"], pa["so"], pa["in"], C.byref(prm), pa["out"],
                                               a_mm, a_ms, a_mc)
    assert rc == 0, rc


for _ in range(warm):
    dev_call()
assert lib.ccicp_multi_sync(m) == 0
t0 = time.perf_counter()
for _ in range(steps):
    dev_call()
assert lib.ccicp_multi_sync(m) == 0
dev_ms = (time.perf_counter() - t0) * 1e3 / steps
same = all(np.array_equal(keep[g][5].cpu().numpy().view(capi.RESULT_DTYPE)["R"], ref["R"]) for g in range(G))
li = capi.LaunchInfo()
lib.ccicp_get_launch_info(lib.ccicp_multi_ctx(m, 0), C.byref(li))
print(json.dumps({"api": "ccicp_register_batch_multi / ccicp_register_batch_multi_device (single process, C ABI)", "devices": devs,
                  "pairs_total": n, "pairs_per_device": P, "steps": steps,
                  "host_path": {"ms_per_call": host_ms, "registrations_per_s": n / (host_ms * 1e-3)},
                  "device_path": {"ms_per_call": dev_ms, "registrations_per_s": n / (dev_ms * 1e-3),
                                  "gather": {1: "nccl", 2: "peer copies"}.get(lib.ccicp_multi_gather_kind(m)),
                                  "every_device_holds_all_records": bool(same)},
                  "slot0_kernel_ms": float(li.last_kernel_ms)}))
lib.ccicp_multi_destroy(m)
