"""TEST INFRASTRUCTURE (oracle): CPU restatement of the ROS tf conversions the reference's facade uses around the
ICP call (ccicp2d/src/icpTools.cpp:168-176 tf::getYaw; :203-209 getEulerYPR + createQuaternionFromRPY).

tf is a third-party dependency of the reference that is neither vendored nor version-pinned
(ccicp2d/package.xml:10-28), so this follows the published LinearMath algorithms; parity for this part is
"unpinned" against the reference itself and is additionally cross-checked against scipy's Rotation in the tests.
Only tests/, __graft_entry__.smoke() and bench.py's CPU leg may import this module."""
import math

import numpy as np


def euler_ypr(q):
    """tf::Matrix3x3(q).getEulerYPR(yaw, pitch, roll, 1) for q = (x, y, z, w), after quaternionMsgToTF."""
    x, y, z, w = (float(v) for v in q)
    n2 = x * x + y * y + z * z + w * w
    if abs(n2 - 1.0) > float(np.float32(0.1)):
        n = math.sqrt(n2)
        x, y, z, w = x * (1.0 / n), y * (1.0 / n), z * (1.0 / n), w * (1.0 / n)
        n2 = x * x + y * y + z * z + w * w
    s = 2.0 / n2
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    m = [[1.0 - (yy + zz), xy - wz, xz + wy],
         [xy + wz, 1.0 - (xx + zz), yz - wx],
         [xz - wy, yz + wx, 1.0 - (xx + yy)]]
    if abs(m[2][0]) >= 1.0:
        yaw = 0.0
        roll = math.atan2(m[2][1], m[2][2])
        pitch = math.pi / 2.0 if m[2][0] < 0 else -math.pi / 2.0
    else:
        pitch = -math.asin(m[2][0])
        roll = math.atan2(m[2][1] / math.cos(pitch), m[2][2] / math.cos(pitch))
        yaw = math.atan2(m[1][0] / math.cos(pitch), m[0][0] / math.cos(pitch))
    return yaw, pitch, roll


def quaternion_from_rpy(roll, pitch, yaw):
    """tf::createQuaternionFromRPY -> (x, y, z, w)."""
    cy, sy = math.cos(yaw * 0.5), math.sin(yaw * 0.5)
    cp, sp = math.cos(pitch * 0.5), math.sin(pitch * 0.5)
    cr, sr = math.cos(roll * 0.5), math.sin(roll * 0.5)
    return (sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy,
            cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy)


def pose_to_Rt(pose):
    """pose = (x, y, z, qx, qy, qz, qw) -> [r00, r01, r10, r11, t0, t1] (icpTools.cpp:168-176)."""
    th = euler_ypr(pose[3:7])[0]
    return np.array([math.cos(th), -math.sin(th), math.sin(th), math.cos(th), pose[0], pose[1]], np.float64)


def Rt_to_pose(Rt, init_pose):
    """icpTools.cpp:195-212."""
    yaw = math.atan2(Rt[2], Rt[0])
    _, pitch, roll = euler_ypr(init_pose[3:7])
    q = quaternion_from_rpy(roll, pitch, yaw)
    return np.array([Rt[4], Rt[5], init_pose[2], q[0], q[1], q[2], q[3]], np.float64)


SENTINEL = np.array([0, 0, 0, 0, 0, 0, 9999.0])


def do_icp_match(refbind, pair, init_pose, max_iter=20, min_delta=1e-6, indist=5.0):
    """CCICP::doICPMatch(model_GA, model_NGA, scene_GA, scene_NGA, initPose) with the compiled reference as the ICP."""
    mg, mn, sg, sn = pair
    if len(sg) + len(sn) < 5:
        return SENTINEL.copy()
    r = refbind.fit(mg, mn, sg, sn, pose_to_Rt(init_pose), max_iter=max_iter, min_delta=min_delta, indist=indist)
    return Rt_to_pose(r["Rt"], init_pose)
