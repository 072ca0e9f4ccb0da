#!/usr/bin/env python
"""Timing of the BASELINE.json configs that are not the bench.py line (C3 sample, C4, C5) on one GPU.
Reports the ICP kernel's own device time (CUDA events inside the library) and the host-call wall time.
usage: python profiles/bench_configs.py [c3] [c4] [c5] [--pairs N]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slam_b200 import capi, scangen  # noqa: E402


def run(ctx, name, b, prm, n_seeds=1, init=None, reps=3):
    init = b.init_Rt() if init is None else init
    out = None
    k_ms, w_ms = [], []
    for _ in range(reps + 1):
        t0 = time.perf_counter()
        out = ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, init, prm, n_seeds=n_seeds)
        w_ms.append((time.perf_counter() - t0) * 1e3)
        k_ms.append(ctx.launch_info().last_kernel_ms)
    li = ctx.launch_info()
    k, w = float(np.min(k_ms[1:])), float(np.min(w_ms[1:]))
    n_prob = b.n_pairs * n_seeds
    print(json.dumps({"config": name, "pairs": b.n_pairs, "seeds": n_seeds, "kernel_ms": k, "host_call_ms": w,
                      "problems_per_s_kernel": n_prob / (k * 1e-3), "problems_per_s_host_call": n_prob / (w * 1e-3),
                      "variant": {0: "on-chip", 1: "l2", 4: "cluster"}.get(int(li.kernel_variant), "?"), "grid": li.grid, "block": li.block,
                      "regs": li.regs_per_thread, "smem": li.smem_bytes, "n_launches": li.n_launches,
                      "mean_num_corr": float(out["num_corr"].mean()), "mean_iters": float(out["iters"].mean()),
                      "tie_replay_pairs": int((out["status"] & capi.ST_TIE_REPLAY != 0).sum())}), flush=True)
    return out


def main():
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c3", "c4", "c5"]
    pairs = int(sys.argv[sys.argv.index("--pairs") + 1]) if "--pairs" in sys.argv else 0
    ctx = capi.Context(0)
    prm = capi.fixed_iter_params(30)
    if "c2" in which:
        run(ctx, "C2 odometry 1081-beam", scangen.generate(1000, pairs or 4096, config=1), prm)
    if "c2x" in which:      # bit-identical mode (sequential sums + NR SVD)
        b2 = scangen.generate(1000, pairs or 4096, config=1, pinned=capi.pinned_array)
        run(ctx, "C2 default mode (pinned)", b2, prm)
        run(ctx, "C2 CCICP_F_EXACT_SUMS (pinned)", b2, capi.fixed_iter_params(30, flags=capi.F_EXACT_SUMS))
    if "c3" in which:
        run(ctx, "C3 loop-closure 1081-beam (pinned host buffers, chunked pipeline)",
            scangen.generate(10**6, pairs or 16384, config=3, pinned=capi.pinned_array), prm)
    if "c4" in which:
        run(ctx, "C4 dense 32768-point", scangen.generate(5000, pairs or 148, config=4), prm, reps=2)
    if "c5" in which:
        P, S = pairs or 256, 256
        b = scangen.generate(2000, P, config=1)
        init = np.concatenate([scangen.pose_to_Rt(scangen.hypotheses(int(s), S)) for s in b.seeds], 0)
        prm5 = capi.fixed_iter_params(30, flags=capi.F_BEST_OF_SEEDS)
        run(ctx, "C5 256 seeds best-of", b, prm5, n_seeds=S, init=init, reps=2)
    if "lum" in which:
        bench_lum(ctx, pairs or 592)
    if "lumbig" in which:      # the size class graph_slam really has: voxel-filtered Velodyne clouds
        bench_lum(ctx, pairs or 64, pts=30000)
    if "classify" in which:
        bench_classify(ctx, pairs or 256)
    ctx.close()


def bench_lum(ctx, n_edges, pts=3000):
    """batched computeEdgeInformationLUM; FP32-issue bound (n x m exact distance sweep)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_lum_gpu import cloud_pair
    from oracle import lum_port
    base = [cloud_pair(k, pts) for k in range(4)]
    src = [base[k % 4][0] for k in range(n_edges)]; tgt = [base[k % 4][1] for k in range(n_edges)]
    ms = []
    for _ in range(4):
        info, num, st = ctx.edge_information_lum(src, tgt, 0.75)
        ms.append(ctx.launch_info().last_kernel_ms)
    k = float(np.min(ms[1:]))
    evals = sum(len(a) * len(b) for a, b in zip(src, tgt))
    t0 = time.perf_counter(); lum_port.edge_information(base[0][0][:600], base[0][1], 0.75); cpu = time.perf_counter() - t0
    print(json.dumps({"config": "computeEdgeInformationLUM", "edges": n_edges, "points_per_cloud": pts, "kernel_ms": k,
                      "edges_per_s": n_edges / (k * 1e-3), "distance_evals_per_s": evals / (k * 1e-3),
                      "fp32_ops_per_s": 8 * evals / (k * 1e-3), "fp32_unfused_peak_ops_per_s": 148 * 128 * 1.965e9,
                      "frac_of_fp32_issue_peak": 8 * evals / (k * 1e-3) / (148 * 128 * 1.965e9),
                      "cpu_port_edges_per_s_est": 1.0 / (cpu * len(base[0][0]) / 600.0), "mean_num_corr": float(num.mean())}), flush=True)


def bench_classify(ctx, n_clouds, pts=70000):
    """batched classifyPoints, device-resident, CUDA-event timed; numpy oracle timed beside it."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_classify_cpu import synthetic_cloud
    from oracle import classify_port as cp
    base = [synthetic_cloud(k, pts) for k in range(8)]
    clouds = [base[k % 8] for k in range(n_clouds)]
    xyz = np.concatenate(clouds, 0)
    off = np.cumsum([0] + [len(c) for c in clouds]).astype(np.int32)
    n = len(xyz)
    dev = torch.device("cuda", 0)
    d_xyz = torch.from_numpy(xyz).to(dev); d_off = torch.from_numpy(off).to(dev)
    o_xyz = torch.empty_like(d_xyz); o_ga = torch.empty(n, dtype=torch.uint8, device=dev)
    o_src = torch.empty(n, dtype=torch.int32, device=dev); o_cnt = torch.empty(n_clouds, dtype=torch.int32, device=dev)
    st = torch.cuda.Stream(device=dev)
    ms = []
    for _ in range(6):
        ctx.classify_points_device(n_clouds, d_xyz.data_ptr(), d_off.data_ptr(), n, o_xyz.data_ptr(), o_ga.data_ptr(),
                                   o_src.data_ptr(), o_cnt.data_ptr(), st.cuda_stream)
        ctx.sync(st.cuda_stream)
        ms.append(ctx.launch_info().last_kernel_ms)
    k = float(np.min(ms[1:]))
    t0 = time.perf_counter()
    for c in base[:4]:
        cp.classify(c)
    cpu_pts_s = 4 * pts / (time.perf_counter() - t0)
    alg = n * (12 + 12 + 1 + 4) + n_clouds * 8            # xyz in; xyz, flag, source index out; offsets + counts
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    print(json.dumps({"config": "classifyPoints", "clouds": n_clouds, "points": n, "ms": k, "points_per_s": n / (k * 1e-3),
                      "clouds_per_s": n_clouds / (k * 1e-3), "n_launches": ctx.launch_info().n_launches,
                      "roofline": {"bound": "hbm", "algorithmic_bytes": alg, "achieved": alg / (k * 1e-3) / 1e9, "peak": peak,
                                   "unit": "GB/s", "frac": alg / (k * 1e-3) / 1e9 / peak},
                      "cpu_baseline": {"value": cpu_pts_s, "unit": "points/s", "cores": 1, "kind": "port (numpy restatement)"},
                      "kept_mean": float(o_cnt.float().mean().item())}), flush=True)


if __name__ == "__main__":
    main()
