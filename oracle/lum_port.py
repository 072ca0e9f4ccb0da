"""TEST INFRASTRUCTURE (oracle): CPU restatement of graph_slam's computeEdgeInformationLUM
(graph_slam/src/graphSlamTools.cpp:108-214) and of the edge rejection rule (:332-362, thresholds
graphSlamTools.h:35-36).

PARITY UNPINNED: the reference function needs PCL (KdTreeFLANN) and Eigen, which are not in this image; it cannot
be compiled here and has no test vectors.  The float / double mix of the reference is kept (Vector3f averages and
differences, float products added into double sums, a float s^2); the nearest neighbour is the exact float
arg-min (scipy cKDTree on float64 copies of the float32 points, re-evaluated in float32), ties -> lowest index.
Only tests/, __graft_entry__.smoke() and bench.py's CPU leg may import this module."""
import numpy as np

f32 = np.float32


def edge_information(source_trans, target, max_corr_distance=0.75):
    s = np.asarray(source_trans, f32).reshape(-1, 3)
    t = np.asarray(target, f32).reshape(-1, 3)
    aver, diff = [], []
    if len(t):
        for i in range(len(s)):
            d = s[i][None, :] - t                                  # float32
            d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
            j = int(np.argmin(d2))                                  # first minimum
            if float(d2[j]) < max_corr_distance * max_corr_distance:
                aver.append(f32(0.5) * (s[i] + t[j]))
                diff.append(s[i] - t[j])
    n = len(aver)
    MM = np.zeros((6, 6)); MZ = np.zeros(6)
    for a, d in zip(aver, diff):                                    # :150-171, float products
        MM[0, 4] -= a[1]; MM[0, 5] += a[2]; MM[1, 3] -= a[2]; MM[1, 4] += a[0]; MM[2, 3] += a[1]; MM[2, 5] -= a[0]
        MM[3, 4] -= a[0] * a[2]; MM[3, 5] -= a[0] * a[1]; MM[4, 5] -= a[1] * a[2]
        MM[3, 3] += a[1] * a[1] + a[2] * a[2]; MM[4, 4] += a[0] * a[0] + a[1] * a[1]; MM[5, 5] += a[0] * a[0] + a[2] * a[2]
        MZ[0] += d[0]; MZ[1] += d[1]; MZ[2] += d[2]
        MZ[3] += a[1] * d[2] - a[2] * d[1]; MZ[4] += a[0] * d[1] - a[1] * d[0]; MZ[5] += a[2] * d[0] - a[0] * d[2]
    MM[0, 0] = MM[1, 1] = MM[2, 2] = float(f32(n))
    for (i, j) in [(4, 0), (5, 0), (3, 1), (4, 1), (3, 2), (5, 2), (4, 3), (5, 3), (5, 4)]:
        MM[i, j] = MM[j, i]
    try:
        D = np.linalg.inv(MM) @ MZ
    except np.linalg.LinAlgError:
        return np.eye(6), n, 1
    ss = f32(0.0)
    for a, d in zip(aver, diff):                                    # :187-193
        e0 = float(d[0]) - (D[0] + float(a[2]) * D[5] - float(a[1]) * D[4])
        e1 = float(d[1]) - (D[1] + float(a[0]) * D[4] - float(a[2]) * D[3])
        e2 = float(d[2]) - (D[2] + float(a[1]) * D[3] - float(a[0]) * D[5])
        ss = f32(ss + f32(e0 * e0 + e1 * e1 + e2 * e2))
    if ss < 0.0000000000001 or not np.isfinite(ss):                 # :196-201
        return np.eye(6), n, 1
    return MM * float(f32(1.0) / ss), n, 0


def edge_is_rejected(init_xyyaw, edge_xyyaw, dist_thresh=10.0, rot_thresh=0.2):
    """graphSlamTools.cpp:332-362: |dx| > 10 m or |dy| > 10 m or wrapped |dyaw| > 0.2 rad -> failed ICP match."""
    xd, yd = abs(init_xyyaw[0] - edge_xyyaw[0]), abs(init_xyyaw[1] - edge_xyyaw[1])
    th = abs(init_xyyaw[2] - edge_xyyaw[2])
    if th > 2 * np.pi:
        th = th - 2 * np.pi
    elif th > np.pi:
        th = 2 * np.pi - th
    return bool(xd > dist_thresh or yd > dist_thresh or th > rot_thresh)
