#!/usr/bin/env python
"""bench.py -- scan-pair registrations/sec of the ccicp2d hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--pairs P]

A "step" is one pass of the hot path over one batch of synthetic scan pairs: config C2 of
BASELINE.json (4096 odometry pairs, 1081 beams each, 30 fixed iterations, indist 5).  Under
torchrun every rank registers its own shard of pairs (weak scaling: 4096 pairs per GPU, no
data-path collective) and the records are all-gathered over NCCL inside the timed region.

  value    whole-job registrations/s with the inputs already resident in HBM
  e2e      the same metric through ccicp_register_batch() with pinned HOST buffers: H2D of the
           scans + init poses and D2H of the result records inside the timed region
  roofline algorithmic HBM bytes of the ICP kernel / its CUDA-event duration vs the measured copy peak
  cpu_baseline  the reference's own CPU code (oracle/_ref: unmodified sources; else the C port)
           timed on this box's host cores on a bounded sample of the same pairs

--impl reference times only that CPU arm and prints the same JSON line with "impl": "reference".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "scan-pair registrations/sec (1081-beam, fixed 30 iters)"
UNIT = "registrations/s"
PAIRS_PER_GPU = 4096
ITERS = 30
INDIST = 5.0
SEED0 = 1000                      # seeds 1000...
C3_SEED0 = 1000000                # seeds 10^6..., offsets +-2 m / +-0.3 rad, init = truth + noise
C3_TOTAL = 262144                 # loop-closure candidates of the whole job (strong scaling: C3_TOTAL / N per GPU)


def generate_chunked(seed_lo, n_pairs, config, pinned, chunk=16384):
    """scangen.generate in chunks (its dense [pairs, beams, 2] temporaries stay small), packed into one pinned CSR."""
    from slam_b200 import scangen
    if n_pairs <= chunk:
        return scangen.generate(seed_lo, n_pairs, config=config, pinned=pinned)
    parts = [scangen.generate(seed_lo + lo, min(chunk, n_pairs - lo), config=config) for lo in range(0, n_pairs, chunk)]
    def cat(xy_name, off_name):
        xy = pinned((sum(len(getattr(b, xy_name)) for b in parts), 2), np.float64)
        off, o = [np.zeros(1, np.int32)], 0
        for b in parts:
            a = getattr(b, xy_name)
            xy[o:o + len(a)] = a
            off.append(getattr(b, off_name)[1:] + np.int32(o))
            o += len(a)
        return xy, np.concatenate(off).astype(np.int32)
    mxy, moff = cat("model_xy", "model_off")
    sxy, soff = cat("scene_xy", "scene_off")
    return scangen.Batch(mxy, moff, sxy, soff, np.concatenate([b.truth for b in parts]),
                         np.concatenate([b.init for b in parts]), np.concatenate([b.seeds for b in parts]))


def algorithmic_bytes(batch, result_bytes=160):
    """inputs as the API delivers them (FP64 xy) read once + offsets + init + record."""
    n_pairs = batch.n_pairs
    pts = len(batch.model_xy) + len(batch.scene_xy)
    return 16 * pts + n_pairs * (16 + 48 + result_bytes)


# ---------------------------------------------------------------- CPU arm ----
def _cpu_worker(args):
    kind, pairs, inits, reps = args
    from oracle import refbind
    t0 = time.perf_counter()
    n = 0
    for _ in range(reps):
        for (mg, mn, sg, sn), Rt0 in zip(pairs, inits):
            refbind.fit(mg, mn, sg, sn, Rt0, max_iter=ITERS, min_delta=-2.0, indist=INDIST, kind=kind)
            n += 1
    return n, time.perf_counter() - t0


def cpu_reference_rate(batch, cores, pairs_per_core, reps=1):
    """Registrations/s of the CPU reference on `cores` host cores: one PROCESS per core (the reference's
    matrix.cpp:26-31 file-static scratch makes it thread-unsafe), each registering its own pairs
    (kd-tree build + 30 fitSteps per pair, like icpTools.cpp:187-188)."""
    import multiprocessing as mp
    from oracle import refbind
    kind = refbind.best_kind()
    refbind.load(kind)
    inits = batch.init_Rt()
    jobs = []
    for c in range(cores):
        idx = [(c * pairs_per_core + j) % batch.n_pairs for j in range(pairs_per_core)]
        jobs.append((kind, [tuple(a.copy() for a in batch.pair(p)) for p in idx], [inits[p].copy() for p in idx], reps))
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_worker, [(kind, jobs[0][1][:1], jobs[0][2][:1], 1)] * cores)     # warm the workers
        t0 = time.perf_counter()
        out = pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
    n = sum(o[0] for o in out)
    return dict(value=n / wall, unit=UNIT, cores=cores, kind="reference" if kind == "reference" else "port",
                sample="%d pairs x %d rep(s) on each of %d processes (first pairs of the same batch), %.1f s wall"
                       % (pairs_per_core, reps, cores, wall)), wall


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ------------------------------------------------------------- clocks --------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------- GPU arm -------
def bind_near_gpu(local_rank):
    """Pins this rank to the CPUs that are local to its GPU (sysfs local_cpulist of the PCI device) BEFORE the pinned
    host buffers are allocated and filled, so that their pages land on the GPU's NUMA node (first touch).  With 8 ranks
    copying at once the host side is the bottleneck of the end-to-end leg.  Returns what it did, for the JSON line."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = open("/sys/bus/pci/devices/%s/numa_node" % bus).read().strip()
        txt = open("/sys/bus/pci/devices/%s/local_cpulist" % bus).read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-"); cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        want = cpus & os.sched_getaffinity(0)
        if not want:
            return {"numa_node": node, "bound": False, "why": "no local CPU in this process's affinity mask"}
        os.sched_setaffinity(0, want)
        return {"numa_node": node, "bound": True, "cpus": len(want)}
    except Exception as e:       # no sysfs entry (VM), old torch: run unbound
        return {"bound": False, "why": str(e)[:80]}


def load_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """dram bytes per launch of icp_pair_kernel from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return None


def issue_roofline(traffic, k_ms, sms, clocks):
    if not traffic or not traffic.get("warp_instr_per_launch"):
        return None
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    peak = sms * 4 * mhz * 1e6                       # 4 schedulers per SM, 1 warp instruction per cycle each
    ach = traffic["warp_instr_per_launch"] / (k_ms * 1e-3)
    return {"bound": "issue", "achieved": ach / 1e12, "peak": peak / 1e12, "unit": "T warp-instr/s", "frac": ach / peak,
            "source": traffic.get("capture")}


def gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from slam_b200 import capi, scangen, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    numa = bind_near_gpu(local_rank) if args.numa_bind else {"bound": False, "why": "--no-numa-bind"}
    c3 = args.workload == "c3"
    if c3:       # the whole loop-closure batch is split over the ranks (strong scaling)
        seed_lo, n_pairs = sharding.shard_seeds(C3_SEED0, args.total_pairs, rank, world)
        assert n_pairs * world == args.total_pairs, "--total-pairs must divide by the number of GPUs"
    else:        # weak scaling: args.pairs per GPU
        n_pairs = args.pairs
        seed_lo, n_local = sharding.shard_seeds(SEED0, n_pairs * world, rank, world)     # contiguous pair ranges
        assert n_local == n_pairs
    batch = generate_chunked(seed_lo, n_pairs, 3 if c3 else 1, capi.pinned_array)
    init_h = capi.pinned_array((n_pairs, 6), np.float64)
    init_h[:] = batch.init_Rt()
    out_h = capi.pinned_array((n_pairs,), capi.RESULT_DTYPE)
    alg_bytes = algorithmic_bytes(batch)
    prm = capi.fixed_iter_params(ITERS, indist=INDIST)
    ctx = capi.Context(local_rank)
    cells_hint = capi.max_cells(batch.model_off)       # what the host-buffer call computes itself from the offsets

    # device-resident copies (the "value" leg starts with inputs in HBM)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
    d_mxy, d_sxy = t(batch.model_xy, np.float64), t(batch.scene_xy, np.float64)
    d_moff, d_soff = t(batch.model_off, np.int32), t(batch.scene_off, np.int32)
    d_init = t(init_h, np.float64)
    d_out = torch.zeros(n_pairs * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)

    def step_device():
        ctx.register_batch_device(n_pairs, 1, d_mxy.data_ptr(), d_moff.data_ptr(), d_sxy.data_ptr(), d_soff.data_ptr(),
                                  d_init.data_ptr(), prm, d_out.data_ptr(), 0, batch.max_model, batch.max_scene,
                                  stream.cuda_stream, max_cells_hint=cells_hint)
        if world > 1:      # the path's only exchange: gather the per-pair records
            return sharding.gather_records(d_out, world, n_pairs * world)
        return d_out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step_device()
        barrier()
        if rank == 0:
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kernel_ms = []
        ev0.record(stream)
        for _ in range(args.steps):
            step_device()
            stream.synchronize()          # per-step so the library's own kernel events can be read back
            kernel_ms.append(ctx.launch_info().last_kernel_ms)
        ev1.record(stream)
        barrier()
        dev_ms = ev0.elapsed_time(ev1) / args.steps
        li = ctx.launch_info()

        # e2e: the public host-buffer call, pinned inputs, H2D + D2H inside the timed region
        for _ in range(max(1, min(args.warmup, 3))):
            ctx.register_batch(batch.model_xy, batch.model_off, batch.scene_xy, batch.scene_off, init_h, prm, out=out_h)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ctx.register_batch(batch.model_xy, batch.model_off, batch.scene_xy, batch.scene_off, init_h, prm, out=out_h)
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        e2e_launches = int(ctx.launch_info().n_launches)      # chunk kernels of one host-buffer call
        barrier()
    clocks = sampler.stop() if rank == 0 else None

    # sanity: the device leg and the host leg computed the same records
    dev_rec = d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    same = bool((dev_rec["num_corr"] == out_h["num_corr"]).all() and np.array_equal(dev_rec["R"], out_h["R"]))
    ok_frac = float((out_h["status"] & ~np.uint32(capi.ST_TIE_REPLAY | capi.ST_CONVERGED) == 0).mean())

    tm = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = tm.tolist()
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return None

    peak, peak_src = load_peak()
    k_ms = float(np.mean(kernel_ms))
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic = load_traffic()
    if traffic:      # the committed capture is one launch of `pairs_per_launch` pairs; both figures are linear in the pairs
        f = n_pairs / float(traffic.get("pairs_per_launch", PAIRS_PER_GPU))
        for k in ("dram_bytes_per_launch", "warp_instr_per_launch"):
            if traffic.get(k):
                traffic[k] = traffic[k] * f
    h2d = int(batch.model_xy.nbytes + batch.scene_xy.nbytes + batch.model_off.nbytes + batch.scene_off.nbytes + init_h.nbytes)
    d2h = int(out_h.nbytes)
    line = {
        "metric": METRIC, "value": world * n_pairs / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms, "higher_is_better": True,
        "scaling": "strong" if c3 else "weak", "vs_baseline": None,
        "dtype": "f32 distances / f64 transform+sums / int32 indices",
        "data": "synthetic (seeded 1081-beam ray-cast rooms, slam_b200/csrc/scangen.c)",
        "config": {"workload": ("C3: %d loop-closure scan pairs in total (%d per GPU), offsets +-2 m / +-0.3 rad, init = truth + "
                                "N(0.3 m, 0.05 rad), 1081 beams, two classes, %d fixed iterations, indist %g"
                                % (n_pairs * world, n_pairs, ITERS, INDIST)) if c3 else
                               "C2: %d odometry scan pairs per GPU, 1081 beams, two classes, %d fixed iterations, indist %g"
                               % (n_pairs, ITERS, INDIST),
                   "pairs_per_gpu": n_pairs, "iters": ITERS, "parallelism": "pairs sharded over %d GPU(s)" % world,
                   "l2": "inputs %.0f MB per GPU > 126 MB L2, no flush" % ((alg_bytes) / 1e6),
                   "collective": "nccl all_gather of 160 B records" if world > 1 else "none"},
        "e2e": {"value": world * n_pairs / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                "d2h_bytes_per_step": d2h * world, "ms_per_step": e2e_ms,
                "api": "ccicp_register_batch (C ABI, pinned host buffers)", "host_numa": numa},
        # kernels of this library launched inside the two timed regions (device leg + host-buffer leg), all ranks
        "gpu_launches": (int(li.n_launches) + e2e_launches) * args.steps * world,
        "gpu_launches_detail": {"device_leg_per_step": int(li.n_launches), "e2e_leg_per_step": e2e_launches},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic.get("dram_bytes_per_launch") if traffic else None,
                     "kernel": "icp_pair_kernel", "kernel_ms": k_ms, "algorithmic_bytes_per_launch": alg_bytes,
                     "peak_source": peak_src,
                     "note": "on-chip kernel: HBM is not the binding resource (SURVEY.md 8(d)); see issue_slot_util in profiles/"},
        # secondary: the resource that actually binds this on-chip kernel is instruction issue;
        # warp instructions per launch come from the committed ncu capture of the same command, time is live
        "roofline_issue": issue_roofline(traffic, k_ms, li.sm_count, clocks),
        "kernel": {"grid": li.grid, "block": li.block, "smem_bytes": li.smem_bytes, "regs": li.regs_per_thread,
                   "ctas_per_sm": li.blocks_per_sm, "sms": li.sm_count},
        "clocks": clocks,
        "check": {"device_leg_equals_host_leg": same, "pairs_status_ok_frac": ok_frac},
    }
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs", type=int, default=PAIRS_PER_GPU, help="scan pairs per GPU per step")
    ap.add_argument("--workload", default="c2", choices=["c2", "c3"],
                    help="c2 (default, the metric's configuration): --pairs odometry pairs per GPU, weak scaling; "
                         "c3: --total-pairs loop-closure candidates split over the GPUs, records gathered over NCCL")
    ap.add_argument("--total-pairs", type=int, default=C3_TOTAL, help="c3 only: pairs of the whole job")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa-bind", dest="numa_bind", action="store_false",
                    help="do not pin the rank to its GPU's local CPUs before allocating the pinned host buffers")
    ap.add_argument("--cpu-pairs-per-core", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return 0
        from slam_b200 import scangen
        cores = host_cores()
        c3 = args.workload == "c3"
        batch = scangen.generate(C3_SEED0 if c3 else SEED0, min(args.pairs, max(64, cores * 8)), config=3 if c3 else 1)
        ppc = args.cpu_pairs_per_core or 48           # ~12 ms each -> ~0.6 s per core per step
        for _ in range(args.warmup):
            cpu_reference_rate(batch, cores, 2)
        vals, walls = [], []
        for _ in range(args.steps):
            cb, wall = cpu_reference_rate(batch, cores, ppc)
            vals.append(cb["value"]); walls.append(wall)
        v = float(np.mean(vals))
        cb["value"] = v
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(walls)) * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32/f64 (CPU)",
                "data": "synthetic (same generator and seeds as the GPU arm)",
                "config": {"workload": "%s sample: %d pairs per core per step, 1081 beams, %d fixed iterations, indist %g"
                                       % (args.workload.upper(), ppc, ITERS, INDIST), "iters": ITERS},
                "cpu_baseline": cb,
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return 0

    cpu_line = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before any CUDA initialisation: the workers are forked
        from slam_b200 import scangen
        cores = host_cores()
        c3 = args.workload == "c3"
        sample = scangen.generate(C3_SEED0 if c3 else SEED0, min(args.pairs, 256), config=3 if c3 else 1)
        ppc = args.cpu_pairs_per_core or 1000        # ~12.5 ms each -> ~12 s of CPU work per core
        cpu_line, _ = cpu_reference_rate(sample, cores, ppc)
    line = gpu_arm(args, rank, world, local_rank)
    if line is not None:
        line["cpu_baseline"] = cpu_line
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
