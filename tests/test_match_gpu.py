"""the batched CCICP::doICPMatch(vectors..., initPose) entry point (ccicp_match_batch and the
header-only C++ CCICPBatch on top of it) against the oracle: tf conversions restated in oracle/tf_port.py around the
compiled reference ICP."""
import os
import struct
import subprocess

import numpy as np
import pytest

from slam_b200 import capi, scangen
from oracle import tf_port

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "tests", "host", "batch_demo_b200")


def _init_poses(b, rng):
    """x,y,yaw of the generator's initial guess, plus a random z / roll / pitch that must be carried through."""
    poses = np.zeros((b.n_pairs, 7))
    for p in range(b.n_pairs):
        x, y, yaw = b.init[p]
        roll, pitch = rng.uniform(-0.2, 0.2, 2)
        poses[p, :3] = x, y, rng.uniform(-1, 1)
        poses[p, 3:] = tf_port.quaternion_from_rpy(roll, pitch, yaw)
    return poses


def _check(out, want, what):
    assert np.abs(out[:, :3] - want[:, :3]).max() <= 1e-5, what
    # same rotation: |<q1,q2>| = 1
    dots = np.abs((out[:, 3:] * want[:, 3:]).sum(1))
    ok = want[:, 6] != 9999
    assert np.all(np.arccos(np.clip(dots[ok], -1, 1)) * 2 <= 1e-6), what
    assert np.array_equal(out[~ok], want[~ok]), what


def test_match_batch_against_oracle(ctx, oracle):
    rng = np.random.default_rng(11)
    b = scangen.generate(300, 24, config=3)
    poses = _init_poses(b, rng)
    prm = capi.default_params()                                  # the facade's own settings: 20 iterations, 1e-6, indist 5
    out, rec = ctx.match_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, poses, prm, want_records=True)
    want = np.stack([tf_port.do_icp_match(oracle, b.pair(p), poses[p]) for p in range(b.n_pairs)])
    _check(out, want, "match_batch")
    assert np.array_equal(out[:, 2], poses[:, 2])                # z passes through untouched
    assert (rec["iters"] <= 20).all()


def test_match_batch_sentinel_for_small_scenes(ctx, oracle):
    rng = np.random.default_rng(2)
    m = rng.uniform(-5, 5, (60, 2))
    s = m[:30] + 0.05
    from helpers import csr_from_pairs
    pairs = [(m[:10], m[10:], s[:2], s[2:4]),        # 4 scene points -> sentinel
             (m[:10], m[10:], s[:6], s[6:]),
             (m[:2], m[2:4], s[:1], s[1:3])]         # too few model points AND scene points -> still the sentinel
    mxy, moff, sxy, soff = csr_from_pairs(pairs)
    poses = np.tile([0.1, -0.2, 0.3, 0, 0, np.sin(0.05), np.cos(0.05)], (3, 1))
    out, rec = ctx.match_batch(mxy, moff, sxy, soff, poses, capi.default_params(), want_records=True)
    assert np.array_equal(out[0], tf_port.SENTINEL) and np.array_equal(out[2], tf_port.SENTINEL)
    want1 = tf_port.do_icp_match(oracle, pairs[1], poses[1])
    _check(out[1:2], want1[None], "small scene neighbour")


@pytest.mark.parametrize("mode", ["batch", "single"])
def test_cpp_batch_facade(tmp_path, oracle, mode):
    if not os.path.exists(DEMO):
        pytest.fail("tests/host/batch_demo_b200 missing: run __graft_entry__.build()")
    rng = np.random.default_rng(4)
    b = scangen.generate(77, 6)
    poses = _init_poses(b, rng)
    path = str(tmp_path / "pairs.bin")
    with open(path, "wb") as f:
        f.write(struct.pack("<i", b.n_pairs))
        for p in range(b.n_pairs):
            pr = b.pair(p)
            f.write(struct.pack("<4i", *[len(a) for a in pr]))
            f.write(struct.pack("<7d", *poses[p]))
            for a in pr:
                f.write(np.ascontiguousarray(a, np.float64).tobytes())
    r = subprocess.run([DEMO, path] + (["single"] if mode == "single" else []), stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    got = np.array([ln.split() for ln in r.stdout.strip().splitlines()], float)
    want = np.stack([tf_port.do_icp_match(oracle, b.pair(p), poses[p]) for p in range(b.n_pairs)])
    _check(got[:, :7], want, "C++ facade " + mode)


def test_match_batch_edge_rejection_rule(ctx, oracle):
    """graph_slam's bad-match test (graphSlamTools.cpp:332-362) on the batched caller: flag iff the result moved more
    than reject_dist in x or y or more than reject_rot in (wrapped) yaw away from the initial pose."""
    from oracle import lum_port
    rng = np.random.default_rng(8)
    b = scangen.generate(500, 12, config=3)
    poses = _init_poses(b, rng)
    prm = capi.default_params()
    prm.reject_dist, prm.reject_rot = 0.3, 0.06                       # tight, so that some of the 12 pairs trip it
    out, rec = ctx.match_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, poses, prm, want_records=True)
    flags = []
    for p in range(b.n_pairs):
        yaw0, yaw1 = tf_port.euler_ypr(poses[p, 3:])[0], tf_port.euler_ypr(out[p, 3:])[0]
        flags.append(lum_port.edge_is_rejected((poses[p, 0], poses[p, 1], yaw0), (out[p, 0], out[p, 1], yaw1), 0.3, 0.06))
    got = (rec["status"] & capi.ST_EDGE_REJECTED) != 0
    assert got.tolist() == flags and 0 < sum(flags) < b.n_pairs
