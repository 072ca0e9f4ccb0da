"""TEST INFRASTRUCTURE (oracle): CPU statement of the curvature / normal features and of the curvature-gated
correspondence search (CCICP_F_CURV_GATE, ccicp_scan_features).

PARITY UNPINNED BY CONSTRUCTION: these are north_star extras with no counterpart on the reference's built path.  The
reference's "cc" is the class constraint (ground-adjacent / not, icpPointToPoint.cpp:59-124); its only normal code is the
dead ccicp2d/src/icpPointToPlane.cpp:276-348 (PCA over the 10 nearest neighbours, normal = eigenvector of the smallest
eigenvalue), which is the shape followed here in 2-D.  With the gate off (the default, and every parity run) nothing in
this file is involved.  Only tests/ may import this module.

Definitions (the GPU side is slam_b200/csrc/features.cuh and the kCurv instantiation of the pair kernel):
  feature(i): neighbours = points j of i's class block with |j - i| <= k and |p_j - p_i|^2 <= radius^2 (i included);
              fewer than 3 -> curvature NaN, normal 0.  Scatter matrix about the neighbours' mean, l1 >= l2 its
              eigenvalues: curvature = l2 / (l1 + l2); normal = unit eigenvector of l2 turned towards the origin.
  gated NN:   for a scene point with curvature c_s the candidates are the model points of its class with
              not(|c_m - c_s| > tol) (NaN compares false: featureless points are compatible with everything); among
              them the float arg-min of the reference's distance expression, ties -> lowest model index; then the
              reference's gate, sums, SVD and compose as in fitStep (icpPointToPoint.cpp:33-172).
"""
import math

import numpy as np


def features(xy, off, k=5, radius=0.5):
    """xy float64 [n,2]; off int32 [2*n_clouds+1] (class blocks).  Returns (normal float32 [n,2], curvature float32 [n])."""
    xy = np.asarray(xy, np.float64).reshape(-1, 2)
    n = len(xy)
    nrm = np.zeros((n, 2), np.float32)
    cur = np.full(n, np.nan, np.float32)
    r2 = radius * radius
    for b in range(len(off) - 1):
        b0, b1 = int(off[b]), int(off[b + 1])
        for i in range(b0, b1):
            px, py = float(xy[i, 0]), float(xy[i, 1])
            lo, hi = max(i - k, b0), min(i + k, b1 - 1)
            sx = sy = 0.0
            m = 0
            for j in range(lo, hi + 1):
                dx, dy = float(xy[j, 0]) - px, float(xy[j, 1]) - py
                if dx * dx + dy * dy <= r2:
                    sx += dx; sy += dy; m += 1
            if m < 3:
                continue
            mx, my = sx / m, sy / m
            cxx = cxy = cyy = 0.0
            for j in range(lo, hi + 1):
                dx, dy = float(xy[j, 0]) - px, float(xy[j, 1]) - py
                if dx * dx + dy * dy <= r2:
                    ux, uy = dx - mx, dy - my
                    cxx += ux * ux; cxy += ux * uy; cyy += uy * uy
            tr, hd = cxx + cyy, (cxx - cyy) * 0.5
            disc = math.sqrt(hd * hd + cxy * cxy)
            l2 = max(tr * 0.5 - disc, 0.0)
            cur[i] = np.float32(l2 / tr) if tr > 0.0 else np.float32(0.0)
            ax, ay, bx, by = cxy, l2 - cxx, l2 - cyy, cxy
            na, nb = ax * ax + ay * ay, bx * bx + by * by
            vx, vy = (ax, ay) if na >= nb else (bx, by)
            nv = math.sqrt(vx * vx + vy * vy)
            if not nv > 0.0:
                continue
            vx, vy = vx / nv, vy / nv
            if vx * px + vy * py > 0.0:
                vx, vy = -vx, -vy
            nrm[i] = (vx, vy)
    return nrm, cur


def _nn_gated(model_f, mcur, q, qcur, tol):
    """float32 arg-min of fl(fl(dx*dx) + fl(dy*dy)) over the compatible model points; ties -> lowest index."""
    idx = np.full(len(q), -1, np.int64)
    dis = np.full(len(q), np.inf, np.float32)
    if len(model_f) == 0:
        return idx, dis
    for i in range(len(q)):
        with np.errstate(invalid="ignore"):
            ok = ~(np.abs(mcur - qcur[i]) > np.float32(tol))
        if not ok.any():
            continue
        dx = model_f[:, 0] - q[i, 0]
        dy = model_f[:, 1] - q[i, 1]
        d = (dx * dx) + (dy * dy)                   # float32, un-fused
        d = np.where(ok, d, np.float32(np.inf))
        j = int(np.argmin(d))                       # first minimum = lowest index
        idx[i], dis[i] = j, d[j]
    return idx, dis


def fit_curv(mGA, mNGA, sGA, sNGA, Rt0, max_iter=30, indist=5.0, tol=0.05, k=5, radius=0.5):
    """Fixed-iteration curvature-gated fit.  Returns dict(Rt, corr [iters, ns] (per-class model index or -1), ncorr, Rt_trace)."""
    blocks_m = [np.asarray(mGA, np.float64).reshape(-1, 2), np.asarray(mNGA, np.float64).reshape(-1, 2)]
    blocks_s = [np.asarray(sGA, np.float64).reshape(-1, 2), np.asarray(sNGA, np.float64).reshape(-1, 2)]
    moff = np.cumsum([0] + [len(b) for b in blocks_m])
    soff = np.cumsum([0] + [len(b) for b in blocks_s])
    _, mcur = features(np.concatenate(blocks_m, 0), moff, k, radius)
    _, scur = features(np.concatenate(blocks_s, 0), soff, k, radius)
    Rt = np.array(Rt0, np.float64).reshape(6).copy()
    ns = int(soff[-1])
    corr_all, ncorr, trace = [], [], [Rt.copy()]
    if sum(len(b) for b in blocks_m) < 5 or ns < 5:
        return dict(Rt=Rt, corr=np.zeros((0, ns), np.int64), ncorr=np.zeros(0, int), Rt_trace=np.array(trace))
    for _ in range(max_iter):
        corr = np.full(ns, -1, np.int64)
        pm, pt = [], []
        for c in range(2):
            M, S = blocks_m[c], blocks_s[c]
            if len(M) <= 3 or len(S) == 0:                      # icpPointToPoint.cpp:59,93
                continue
            Mf = M.astype(np.float32)
            q = np.stack([((Rt[0] * S[:, 0] + Rt[1] * S[:, 1]) + Rt[4]).astype(np.float32),
                          ((Rt[2] * S[:, 0] + Rt[3] * S[:, 1]) + Rt[5]).astype(np.float32)], 1)
            idx, dis = _nn_gated(Mf, mcur[moff[c]:moff[c + 1]], q, scur[soff[c]:soff[c + 1]], tol)
            keep = (idx >= 0) & (dis.astype(np.float64) < indist)
            corr[soff[c]:soff[c + 1]] = np.where(keep, idx, -1)
            pm.append(Mf[idx[keep]].astype(np.float64)); pt.append(q[keep].astype(np.float64))
        corr_all.append(corr)
        n = int((corr >= 0).sum())
        ncorr.append(n)
        if n == 0:
            trace.append(Rt.copy())
            continue
        p_m, p_t = np.concatenate(pm, 0), np.concatenate(pt, 0)
        mu_m, mu_t = p_m.sum(0) / n, p_t.sum(0) / n
        H = (p_t - mu_t).T @ (p_m - mu_m)
        U, _, Vt = np.linalg.svd(H)
        R_ = Vt.T @ U.T                                          # no det fix, like the reference (:162)
        t_ = mu_m - R_ @ mu_t
        R = R_ @ Rt[:4].reshape(2, 2)
        t = R_ @ Rt[4:] + t_
        Rt = np.r_[R.reshape(4), t]
        trace.append(Rt.copy())
    return dict(Rt=Rt, corr=np.array(corr_all), ncorr=np.array(ncorr), Rt_trace=np.array(trace))
