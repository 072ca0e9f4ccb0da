"""Shared helpers for the parity tests (test scaffolding)."""
import numpy as np

from slam_b200 import capi, scangen

TOL_T = 1e-5      # metres   (BASELINE.json north_star)
TOL_YAW = 1e-6    # radians


def csr_from_pairs(pairs):
    """pairs: list of (mGA, mNGA, sGA, sNGA) float64 [n,2] arrays -> CSR arrays."""
    mo, so, mx, sx = [0], [0], [], []
    for mg, mn, sg, sn in pairs:
        for arr, off, dst in ((mg, mo, mx), (mn, mo, mx), (sg, so, sx), (sn, so, sx)):
            arr = np.asarray(arr, np.float64).reshape(-1, 2)
            off.append(off[-1] + len(arr))
            dst.append(arr)
    cat = lambda l: np.ascontiguousarray(np.concatenate(l, 0)) if l else np.zeros((0, 2))
    return cat(mx), np.array(mo, np.int32), cat(sx), np.array(so, np.int32)


def yaw_of(Rt):
    Rt = np.asarray(Rt).reshape(-1, 6)
    return np.arctan2(Rt[:, 2], Rt[:, 0])


def assert_pose_close(Rt_gpu, Rt_ref, what=""):
    Rt_gpu = np.asarray(Rt_gpu).reshape(-1, 6)
    Rt_ref = np.asarray(Rt_ref).reshape(-1, 6)
    dt = np.abs(Rt_gpu[:, 4:] - Rt_ref[:, 4:]).max()
    dy = np.abs(np.angle(np.exp(1j * (yaw_of(Rt_gpu) - yaw_of(Rt_ref))))).max()
    dR = np.abs(Rt_gpu[:, :4] - Rt_ref[:, :4]).max()
    assert dt <= TOL_T and dy <= TOL_YAW and dR <= 1e-6, "%s: |dt|=%g m |dyaw|=%g rad |dR|=%g" % (what, dt, dy, dR)
    return dt, dy


def gpu_lockstep_corr(ctx, pair, Rt_trace, indist=5.0, flags=0):
    """Correspondences of ONE fitStep from each R,t of the oracle's trace: the pair is submitted once with
    n_seeds = len(trace) initial transforms and max_iter = 1.  Returns (corr[k, ns], results)."""
    mxy, moff, sxy, soff = csr_from_pairs([pair])
    k = len(Rt_trace)
    prm = capi.default_params(max_iter=1, min_delta=-2.0, indist=indist, flags=flags)
    res, corr = ctx.register_batch(mxy, moff, sxy, soff, np.asarray(Rt_trace), prm, n_seeds=k, want_corr=True)
    return corr, res


# ---- the oracle over many problems, forked over the host cores (the reference is not thread-safe:
# matrix.cpp:26-31 file-static scratch; one PROCESS per core like bench.py's CPU arm) -------------------------
def _oracle_job(job):
    from oracle import refbind
    (pair, Rt0, kw), opts = job
    tr = refbind.trace(*pair, Rt0, want_tie=False, **kw)
    out = dict(iters=tr["iters"], status=tr["status"], Rt=tr["Rt"])
    ok = np.nonzero(tr["ncorr"] > 0)[0]
    if len(ok):                      # numCorr / cModelPoints of the reference: the last fitStep that had correspondences
        k = ok[-1]
        out["num_corr"] = int(tr["ncorr"][k])
        out["score"] = float(tr["dis"][k][tr["corr"][k] >= 0].astype(np.float64).mean())
    else:
        out["num_corr"], out["score"] = 0, float("inf")
    if opts.get("last_corr"):
        out["corr"] = tr["corr"][-1].copy() if tr["iters"] else np.zeros(0, np.int32)
    if opts.get("all_corr"):
        out["corr"] = tr["corr"].copy()
    return out


def oracle_map(jobs, **opts):
    """jobs: list of (pair, Rt0, trace kwargs).  Returns one small dict per job (see _oracle_job)."""
    import multiprocessing as mp
    import os
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    jobs = [(j, opts) for j in jobs]
    if cores <= 1 or len(jobs) < 4:
        return [_oracle_job(j) for j in jobs]
    # fork: the children only run the CPU oracle (never CUDA) and leave through os._exit
    with mp.get_context("fork").Pool(min(cores, 32)) as pool:
        return pool.map(_oracle_job, jobs, chunksize=max(1, len(jobs) // (cores * 8)))
