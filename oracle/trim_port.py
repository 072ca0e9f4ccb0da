"""TEST INFRASTRUCTURE -- CPU statement of the CCICP_F_TRIM extra (trimmed outlier rejection).

PARITY UNPINNED: this is a north_star extra with NO counterpart in the reference (its only outlier test is the fixed
gate `dis < inDist`, ccicp2d/src/icpPointToPoint.cpp:76,111).  Everything else follows the reference's fitStep
(icpPointToPoint.cpp:33-172): FP64 un-fused transform rounded to a float query, the kd-tree's exact float NN with its own
tie rule (answered by the compiled reference / the C port through `refbind.nn`), the strict gate, GA block then NGA
block, classes with <= 3 model points skipped, R_ = V * U^T of the 2x2 cross-covariance without a reflection fix,
fitStep -> -1 on no correspondences.  The addition: of the n gated correspondences of an iteration only those whose
squared distance is <= the k-th smallest, k = ceil(trim_frac * n) clamped to [1, n], take part in the solve.
Only tests/ may import this module."""
import math

import numpy as np

from oracle import refbind


def _svd_rot(H):
    """R_ = V * U^T as Matrix::svd + icpPointToPoint.cpp:160-161 give it (no det fix)."""
    U, _, Vt = np.linalg.svd(H)
    return Vt.T @ U.T


def fit_trimmed(mGA, mNGA, sGA, sNGA, Rt0, max_iter=30, min_delta=-2.0, indist=5.0, trim_frac=1.0):
    """Returns dict(Rt, corr[ns] of the last executed iteration (-1 = none), num_corr, iters)."""
    models = [np.ascontiguousarray(mGA, np.float64).reshape(-1, 2), np.ascontiguousarray(mNGA, np.float64).reshape(-1, 2)]
    scenes = [np.ascontiguousarray(sGA, np.float64).reshape(-1, 2), np.ascontiguousarray(sNGA, np.float64).reshape(-1, 2)]
    Rt = np.array(Rt0, np.float64).reshape(6).copy()
    ns = len(scenes[0]) + len(scenes[1])
    corr = np.full(ns, -1, np.int32)
    num_corr, iters = 0, 0
    for _ in range(max_iter):
        iters += 1
        idx_all = np.full(ns, -1, np.int32)
        dis_all = np.full(ns, np.inf, np.float32)
        q_all = np.zeros((ns, 2), np.float32)
        o = 0
        for M, S in zip(models, scenes):
            n = len(S)
            if n and len(M) > 3:
                qx = ((Rt[0] * S[:, 0] + Rt[1] * S[:, 1]) + Rt[4]).astype(np.float32)
                qy = ((Rt[2] * S[:, 0] + Rt[3] * S[:, 1]) + Rt[5]).astype(np.float32)
                q = np.stack([qx, qy], 1)
                idx, dis = refbind.nn(M, q)
                keep = dis.astype(np.float64) < indist
                idx_all[o:o + n] = np.where(keep, idx, -1)
                dis_all[o:o + n] = np.where(keep, dis, np.inf)
                q_all[o:o + n] = q
            o += n
        gated = idx_all >= 0
        n_g = int(gated.sum())
        if n_g:
            k = min(max(int(math.ceil(trim_frac * n_g)), 1), n_g)
            thr = np.partition(dis_all[gated], k - 1)[k - 1]
            idx_all[gated & (dis_all > thr)] = -1
        corr = idx_all
        keep = idx_all >= 0
        n = int(keep.sum())
        if n == 0:
            if -1.0 < min_delta:
                break
            continue
        num_corr = n
        pm = np.concatenate([models[0], models[1]], 0).astype(np.float32).astype(np.float64)   # model as float (icp.cpp:51-60)
        off = np.r_[np.zeros(len(scenes[0]), np.int64), np.full(len(scenes[1]), len(models[0]), np.int64)]
        P_m = pm[(idx_all + off)[keep]]
        P_t = q_all[keep].astype(np.float64)
        mu_m, mu_t = P_m.mean(0), P_t.mean(0)
        H = (P_t - mu_t).T @ (P_m - mu_m)
        R_ = _svd_rot(H)
        t_ = mu_m - R_ @ mu_t
        R = R_ @ Rt[:4].reshape(2, 2)
        t = R_ @ Rt[4:] + t_
        Rt = np.r_[R.reshape(-1), t]
        delta = max(np.linalg.norm(R_ - np.eye(2)), np.linalg.norm(t_))
        if delta < min_delta:
            break
    return dict(Rt=Rt, corr=corr, num_corr=num_corr, iters=iters)
