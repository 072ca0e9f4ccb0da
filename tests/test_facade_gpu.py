"""Drop-in check of the C++ boundary: ONE source file written like the reference's call site
(tests/host/callsite_demo.cpp ~ ccicp2d/src/icpTools.cpp:168-197) is compiled against (a) the drop-in headers +
libccicp2d.so and (b) the reference's own headers + unmodified sources; both binaries must print the same
registration for the same pair."""
import os
import struct
import subprocess

import numpy as np
import pytest

from slam_b200 import scangen

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
B200 = os.path.join(ROOT, "tests", "host", "callsite_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "callsite_ref")


def _write_pair(path, pair, pose):
    with open(path, "wb") as f:
        f.write(struct.pack("<4i", *[len(a) for a in pair]))
        f.write(struct.pack("<3d", *pose))
        for a in pair:
            f.write(np.ascontiguousarray(a, np.float64).tobytes())


def _run(exe, path, extra):
    r = subprocess.run([exe, path] + extra, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    return np.array(lines[0].split(), float), int(lines[1]), np.array(lines[2].split(), float), r.stderr


@pytest.mark.parametrize("seed,config,extra", [(1, 1, ["30", "-2"]), (21, 1, []), (1000007, 3, ["30", "-2"])])
def test_call_site_compiles_and_agrees(tmp_path, seed, config, extra):
    if not os.path.exists(B200):
        pytest.fail("tests/host/callsite_b200 missing: run __graft_entry__.build()")
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/callsite_ref not built (needs /root/reference at build time)")
    b = scangen.generate(seed, 1, config=config)
    path = str(tmp_path / "pair.bin")
    _write_pair(path, b.pair(0), b.init[0])
    pg, ng, eg, _ = _run(B200, path, extra)
    pr, nr, er, _ = _run(REF, path, extra)
    assert ng == nr
    assert np.abs(pg[:2] - pr[:2]).max() <= 1e-5 and abs(pg[2] - pr[2]) <= 1e-6
    np.testing.assert_allclose(eg, er, rtol=1e-6, atol=1e-6 * np.abs(er).max())


def test_too_few_points_leave_pose_untouched(tmp_path):
    rng = np.random.default_rng(0)
    big = rng.uniform(-5, 5, (40, 2))
    path = str(tmp_path / "few.bin")
    _write_pair(path, (big[:10], big[10:], big[:2], big[2:4]), (0.25, -0.5, 0.125))     # 4 scene points
    pg, ng, _, err = _run(B200, path, ["5", "-2"])
    assert np.allclose(pg, [0.25, -0.5, 0.125], atol=1e-15) and "Total template has 4 points" in err


def test_several_host_threads_share_the_class_surface():
    """SURVEY.md 8(b) threading: four host threads construct IcpPointToPoint objects and call fit() at the same time
    (the reference cannot: matrix.cpp:26-31 file-static scratch); every answer equals the single-thread run."""
    exe = os.path.join(ROOT, "tests", "host", "facade_threads_b200")
    if not os.path.exists(exe):
        pytest.fail("tests/host/facade_threads_b200 missing: run __graft_entry__.build()")
    r = subprocess.run([exe, "4", "6", "5"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "facade_threads OK" in r.stdout, r.stdout
