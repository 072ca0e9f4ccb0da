"""TEST INFRASTRUCTURE (oracle): CPU restatement of CCICP::classifyPoints
(ccicp2d/src/icpTools.cpp:36-103; NUMBINSGA 1200, RESOLUTION 0.5, GRD_ADJ_THRESH 2: icpTools.h:24-26).

PINNED: oracle/Makefile cuts the function (lines 36-103) out of the reference file at build time and compiles it
unmodified against a small pcl::PointCloud shim (oracle/shim_pcl/classify_shim.h) into oracle/_ref/libclassify_ref.so;
tests/test_classify_cpu.py checks this port against it (order, coordinates, flags).  The port stays because it also
returns the source index of every kept point, which the reference does not carry, and as the fallback where the
compiled library is absent.  `classify_literal` follows the reference loop by loop, `classify` is the vectorised form.
Only tests/, __graft_entry__.smoke() and bench.py's CPU leg may import this module."""
import math

import numpy as np

NUMBINSGA, RESOLUTION, GRD_ADJ_THRESH = 1200, 0.5, 2


def classify_literal(xyz):
    """Line-by-line restatement (icpTools.cpp:39-102).  xyz float32 [n,3] -> (xyz_out, ground_adj, src_index)."""
    xyz = np.asarray(xyz, np.float32).reshape(-1, 3)
    bins = {}
    offset = float(NUMBINSGA) * RESOLUTION / 2
    for i in range(len(xyz)):
        cx, cy = float(xyz[i, 0]), float(xyz[i, 1])              # float -> double promotion
        if math.isnan(cx) or math.isnan(cy) or math.isinf(cx) or math.isinf(cy):
            continue
        bx, by = math.floor((cx + offset) / RESOLUTION), math.floor((cy + offset) / RESOLUTION)
        if bx < 0 or bx >= NUMBINSGA or by < 0 or by >= NUMBINSGA:
            continue
        bins.setdefault(bx * NUMBINSGA + by, []).append(i)
    out, ga, src = [], [], []
    for h in sorted(bins):                                        # the i, j double loop visits bins in this order
        i, j = divmod(h, NUMBINSGA)
        if i == 0 or i == NUMBINSGA - 1 or j == 0 or j == NUMBINSGA - 1:
            continue
        ground = 0
        for q in range(i - 1, i + 2):
            for r in range(j - 1, j + 2):
                if q == i and r == j:
                    continue
                if (q * NUMBINSGA + r) not in bins:
                    ground += 1
        for k in bins[h]:
            out.append(xyz[k]); ga.append(1 if ground >= GRD_ADJ_THRESH else 0); src.append(k)
    return (np.array(out, np.float32).reshape(-1, 3), np.array(ga, np.uint8), np.array(src, np.int32))


def classify(xyz):
    """Vectorised restatement: same outputs as classify_literal."""
    xyz = np.asarray(xyz, np.float32).reshape(-1, 3)
    offset = float(NUMBINSGA) * RESOLUTION / 2
    with np.errstate(invalid="ignore"):
        fx = np.floor((xyz[:, 0].astype(np.float64) + offset) / RESOLUTION)
        fy = np.floor((xyz[:, 1].astype(np.float64) + offset) / RESOLUTION)
        ok = (fx >= 0) & (fx < NUMBINSGA) & (fy >= 0) & (fy < NUMBINSGA)
    idx = np.nonzero(ok)[0]
    bx, by = fx[idx].astype(np.int64), fy[idx].astype(np.int64)
    occ = np.zeros((NUMBINSGA, NUMBINSGA), bool)
    occ[bx, by] = True
    inner = (bx > 0) & (bx < NUMBINSGA - 1) & (by > 0) & (by < NUMBINSGA - 1)
    idx, bx, by = idx[inner], bx[inner], by[inner]
    order = np.argsort(bx * NUMBINSGA + by, kind="stable")
    idx, bx, by = idx[order], bx[order], by[order]
    empty = np.zeros(len(idx), np.int32)
    for dq in (-1, 0, 1):
        for dr in (-1, 0, 1):
            if dq or dr:
                empty += ~occ[bx + dq, by + dr]
    return xyz[idx], (empty >= GRD_ADJ_THRESH).astype(np.uint8), idx.astype(np.int32)
