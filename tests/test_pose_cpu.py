"""Host-side pose conversions of ccicp_match_batch (no GPU needed: plain host functions of the C-ABI library)
against the oracle restatement of ROS tf and against scipy's independent implementation."""
import numpy as np
from scipy.spatial.transform import Rotation

from slam_b200 import capi
from oracle import tf_port


def _poses(n=400, seed=3):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((n, 4))
    q[: n // 2] /= np.linalg.norm(q[: n // 2], axis=1, keepdims=True)      # half normalised, half not
    q[n // 2:] *= rng.uniform(0.2, 3.0, (n - n // 2, 1))
    p = np.c_[rng.uniform(-50, 50, (n, 3)), q]
    # hand-made corners: identity, pure yaw, gimbal lock up / down
    extra = [[1, 2, 3, 0, 0, 0, 1], [0, 0, 0, 0, 0, np.sin(0.35), np.cos(0.35)],
             [0, 0, 0, 0, np.sin(np.pi / 4), 0, np.cos(np.pi / 4)], [0, 0, 0, 0, -np.sin(np.pi / 4), 0, np.cos(np.pi / 4)]]
    return np.r_[p, np.array(extra, np.float64)]


def test_pose_to_Rt_matches_oracle_and_scipy():
    for p in _poses():
        Rt = capi.pose_to_Rt(p)
        assert np.array_equal(Rt, tf_port.pose_to_Rt(p))                   # same libm, same expression order
        yaw = Rotation.from_quat(p[3:7] / np.linalg.norm(p[3:7])).as_euler("ZYX")[0]
        m20 = Rotation.from_quat(p[3:7] / np.linalg.norm(p[3:7])).as_matrix()[2, 0]
        if abs(m20) < 1 - 1e-9:                                           # away from gimbal lock the yaw is unique
            assert abs(np.angle(np.exp(1j * (np.arctan2(Rt[2], Rt[0]) - yaw)))) < 1e-9


def test_Rt_to_pose_matches_oracle_and_keeps_roll_pitch():
    rng = np.random.default_rng(5)
    for p in _poses():
        th = rng.uniform(-np.pi, np.pi)
        Rt = np.array([np.cos(th), -np.sin(th), np.sin(th), np.cos(th), rng.uniform(-9, 9), rng.uniform(-9, 9)])
        out = capi.Rt_to_pose(Rt, p)
        assert np.array_equal(out, tf_port.Rt_to_pose(Rt, p))
        assert out[0] == Rt[4] and out[1] == Rt[5] and out[2] == p[2]
        q = p[3:7] / np.linalg.norm(p[3:7])
        if abs(Rotation.from_quat(q).as_matrix()[2, 0]) < 1 - 1e-6:
            y0, p0, r0 = Rotation.from_quat(q).as_euler("ZYX")
            y1, p1, r1 = Rotation.from_quat(out[3:7]).as_euler("ZYX")
            assert abs(p1 - p0) < 1e-8 and abs(np.angle(np.exp(1j * (r1 - r0)))) < 1e-8
            assert abs(np.angle(np.exp(1j * (y1 - th)))) < 1e-8
        assert abs(np.linalg.norm(out[3:7]) - 1) < 1e-12


def test_known_answers():
    # yaw-only quaternion: R = rot(0.7); tf::getYaw of (0,0,sin(.35),cos(.35)) is 0.7
    Rt = capi.pose_to_Rt([4, 5, 6, 0, 0, np.sin(0.35), np.cos(0.35)])
    np.testing.assert_allclose(Rt, [np.cos(0.7), -np.sin(0.7), np.sin(0.7), np.cos(0.7), 4, 5], atol=1e-15)
    # gimbal lock: tf reports yaw = 0
    assert np.array_equal(capi.pose_to_Rt([0, 0, 0, 0, np.sin(np.pi / 4), 0, np.cos(np.pi / 4)])[:4], [1, -0.0, 0, 1])
