"""The multi-device entry points of the C ABI (ccicp_multi_*): a C++ host (tests/host/multi_demo.cu, no Python, no
torch in the process) splits a batch over device slots, on the host-buffer path and on the device-resident path (NCCL
or peer-copy gather), and must reproduce the single-device records exactly.  On a one-GPU box the two slots share
device 0; with >= 2 GPUs the same binary runs over distinct devices and the gather goes over NCCL."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from slam_b200 import capi, scangen

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "tests", "host", "multi_demo_b200")


def _run(devs, pairs):
    assert os.path.exists(DEMO), "tests/host/multi_demo_b200 missing: run __graft_entry__.build()"
    r = subprocess.run([DEMO, devs, str(pairs)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "multi_demo OK" in r.stdout, r.stdout
    return r.stdout


def test_cpp_host_two_slots_on_one_device():
    out = _run("0,0", 97)
    assert "peer copies" in out                    # a device listed twice cannot form an NCCL communicator


def test_cpp_host_three_slots_ragged():
    _run("0,0,0", 50)


def test_cpp_host_distinct_devices_nccl():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU on this box: the NCCL gather needs >= 2 distinct devices")
    out = _run(",".join(str(i) for i in range(min(n, 8))), 1000)
    assert "gather = nccl" in out, out


def test_python_binding_of_the_host_path(ctx):
    """ccicp_register_batch_multi through ctypes, 2 slots, best-of-seeds records included."""
    lib = capi.load()
    vp = C.c_void_p
    lib.ccicp_multi_create.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.ccicp_register_batch_multi.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, C.POINTER(capi.Params), vp, vp]
    lib.ccicp_multi_destroy.argtypes = [vp]
    devs = (C.c_int * 2)(0, 0)
    m = vp()
    assert lib.ccicp_multi_create(devs, 2, C.byref(m)) == capi.OK
    b = scangen.generate(2000, 9)
    n_seeds = 5
    hyp = np.stack([scangen.hypotheses(int(s), n_seeds) for s in b.seeds])
    hyp[:, 1:, :2] *= 0.03
    hyp[:, 1:, 2] = (hyp[:, 1:, 2] - np.pi) * 0.02
    Rt0 = np.ascontiguousarray(scangen.pose_to_Rt(hyp.reshape(-1, 3)))
    P = lambda a: a.ctypes.data_as(vp)
    for flags, n_out in ((0, 9 * n_seeds), (capi.F_BEST_OF_SEEDS, 9)):
        prm = capi.fixed_iter_params(12, flags=flags)
        want = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, Rt0, prm, n_seeds=n_seeds)
        got = np.zeros(n_out, capi.RESULT_DTYPE)
        rc = lib.ccicp_register_batch_multi(m, 9, n_seeds, P(b.model_xy), P(b.model_off), P(b.scene_xy), P(b.scene_off), P(Rt0),
                                            C.byref(prm), P(got), None)
        assert rc == capi.OK
        for f in ("R", "t", "num_corr", "iters", "best_seed", "score"):
            assert np.array_equal(got[f], want[f]), f
    assert lib.ccicp_multi_destroy(m) == capi.OK
