#!/usr/bin/env python
"""bench.py -- scan-pair registrations/sec of the ccicp2d hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--pairs P]

A "step" is one pass of the hot path over one batch of synthetic scan pairs: config C2 of
BASELINE.json (4096 odometry pairs, 1081 beams each, 30 fixed iterations, indist 5).  Under
torchrun every rank registers its own shard of pairs (weak scaling: 4096 pairs per GPU, no
data-path collective) and the records are all-gathered over NCCL inside the timed region.
The same JSON line carries a `configs` object with the other BASELINE.json configs measured the
same way (device-resident + end to end, roofline, cpu_baseline): C4 (1024 pairs x 32768 points)
and C5 (1024 pairs x 256 seeds, best seed on device) at 1 GPU, C3 (262144 loop-closure pairs
split over the ranks, strong scaling) under torchrun.

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

T_START = time.perf_counter()      # the whole run's wall clock is reported as bench_wall_s
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
C4_SEED0, C4_PAIRS = 5000, 1024   # 1024 pairs of 32768-beam dense slices
C5_SEED0, C5_PAIRS, C5_SEEDS = 2000, 1024, 256   # 256 seeds x 1024 pairs, best seed on device


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


def cpu_reference_rate(batch, cores, pairs_per_core, reps=1, inits=None, seeds_per_pair=1, unit=UNIT, what="pairs"):
    """Registrations/s of the CPU reference on `cores` host cores: one PROCESS per core (the reference's
    matrix.cpp:26-31 file-static scratch makes it thread-unsafe), each registering its own problems
    (kd-tree build + 30 fitSteps per problem, like icpTools.cpp:187-188).  inits: [n_pairs * seeds_per_pair, 6]
    initial transforms (default: the batch's own); problem q = (pair q // seeds_per_pair, seed q % seeds_per_pair)."""
    import multiprocessing as mp
    from oracle import refbind
    kind = refbind.best_kind()
    refbind.load(kind)
    inits = batch.init_Rt() if inits is None else inits
    n_prob = batch.n_pairs * seeds_per_pair
    jobs = []
    for c in range(cores):
        idx = [(c * pairs_per_core + j) % n_prob for j in range(pairs_per_core)]
        jobs.append((kind, [tuple(a.copy() for a in batch.pair(q // seeds_per_pair)) for q in idx], [inits[q].copy() for q in idx], reps))
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_worker, [(kind, jobs[0][1][:1], jobs[0][2][:1], 1)] * cores)     # warm the workers
        t0 = time.perf_counter()
        out = pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
    n = sum(o[0] for o in out)
    return dict(value=n / wall, unit=unit, cores=cores, kind="reference" if kind == "reference" else "port",
                sample="%d %s x %d rep(s) on each of %d processes (first %s of the same batch), %.1f s wall"
                       % (pairs_per_core, what, reps, cores, what, wall)), wall


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


def lib_sha256():
    import hashlib
    from slam_b200 import capi
    try:
        return hashlib.sha256(open(capi.LIB_PATH, "rb").read()).hexdigest()
    except Exception:
        return None


def load_traffic():
    """Per-kernel figures of the committed ncu --set full captures (profiles/traffic.json): dram bytes and warp
    instructions per launch.  The file is stamped with the SHA-256 of the library it was captured from; when that is not
    the library loaded now, everything derived from it is reported as stale instead of as a number."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        t = json.load(open(p))
    except Exception:
        return None
    # same binary, or (nvcc builds are not bit-reproducible) a rebuild from the same sources and flags
    from slam_b200.build import source_sha256
    if t.get("lib_sha256") == lib_sha256():
        t["stale"], t["match"] = False, "library"
    elif t.get("src_sha256") and t.get("src_sha256") == source_sha256():
        t["stale"], t["match"] = False, "sources"
    else:
        t["stale"], t["match"] = True, None
    return t


def kernel_traffic(traffic, kernel, units):
    """(dram bytes per launch, warp instructions per launch, capture name) of `kernel` scaled to `units` work units."""
    if not traffic or traffic.get("stale"):
        return None, None, None
    k = (traffic.get("kernels") or {}).get(kernel)
    if not k:
        return None, None, None
    f = units / float(k["units_per_launch"])
    return k["dram_bytes_per_launch"] * f, (k.get("warp_instr_per_launch") or 0) * f or None, k.get("capture")


def issue_roofline(traffic, kernel, units, k_ms, sms, clocks):
    """Issue-slot occupancy: warp instructions of the committed capture / live kernel time vs 4 schedulers x SM clock.
    It says how busy the schedulers are, NOT how much of the work is needed (see roofline_useful)."""
    if traffic and traffic.get("stale"):
        return {"bound": "issue", "frac": None, "stale": True,
                "why": "profiles/traffic.json was captured from a library built from other sources (lib_sha256 and src_sha256 differ)"}
    _, instr, cap = kernel_traffic(traffic, kernel, units)
    if not instr:
        return None
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    peak = sms * 4 * mhz * 1e6                       # 4 schedulers per SM, 1 warp instruction per cycle each
    ach = instr / (k_ms * 1e-3)
    return {"bound": "issue", "achieved": ach / 1e12, "peak": peak / 1e12, "unit": "T warp-instr/s", "frac": ach / peak,
            "source": cap, "stale": False}


def load_fp32_peak():
    """Measured issue peak of un-fused FP32 (profiles/fp32_peak.cu -> profiles/fp32_unfused_peak.json)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "fp32_unfused_peak.json")))
        return float(d["unfused_fp32_lane_ops_per_s"]), "measured (profiles/fp32_unfused_peak.json: FMUL/FADD issue rate, all SMs)"
    except Exception:
        return 148 * 128 * 1.965e9, "nominal 148 SM x 128 lanes x 1.965 GHz (profiles/fp32_unfused_peak.json missing)"


def useful_roofline(counts, launches, k_ms):
    """FP32 utilisation judged on USEFUL operations, not on operations issued.  counts = (distance
    evaluations, queries, accepted pairs) of `launches` diagnostic launches of the same batch; per launch the useful
    lane-operations are  5 per distance evaluation (2 sub, 2 mul, 1 add)  +  8 per query (FP64 transform: 4 mul, 4 add)
    +  16 per accepted pair (shifted sums: 4 sub, 4 add, 4 mul, 4 add)."""
    ev, qu, inl = [c / float(launches) for c in counts]
    ops = 5.0 * ev + 8.0 * qu + 16.0 * inl
    peak, src = load_fp32_peak()
    ach = ops / (k_ms * 1e-3)
    return {"bound": "fp32 (un-fused)", "achieved": ach / 1e12, "peak": peak / 1e12, "unit": "T lane-op/s", "frac": ach / peak,
            "peak_source": src, "distance_evals_per_launch": ev, "queries_per_launch": qu, "accepted_pairs_per_launch": inl,
            "evals_per_query": ev / max(qu, 1.0), "reference_evals_per_query": 11.1,
            "note": "useful ops = 5 x evals + 8 x queries + 16 x accepted pairs (counted by the CCICP_F_DIAG_COUNTS instantiation "
                    "outside the timed region); the reference's kd-tree evaluates 11.1 leaf points per query (SURVEY.md 8(d))"}


class DeviceBatch:
    """One workload resident in HBM (the `value` leg) plus its pinned host copy (the `e2e` leg)."""

    def __init__(self, torch, dev, capi, batch, init_Rt, n_seeds, prm):
        self.capi, self.batch, self.n_seeds, self.prm = capi, batch, n_seeds, prm
        self.n_pairs = batch.n_pairs
        self.best_of = bool(prm.flags & capi.F_BEST_OF_SEEDS) and n_seeds > 1
        self.n_out = self.n_pairs if self.best_of else self.n_pairs * n_seeds
        self.init_h = capi.pinned_array((self.n_pairs * n_seeds, 6), np.float64)
        self.init_h[:] = init_Rt
        self.out_h = capi.pinned_array((self.n_out,), capi.RESULT_DTYPE)
        t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt).reshape(-1)).to(dev)
        self.d = [t(batch.model_xy, np.float64), t(batch.model_off, np.int32), t(batch.scene_xy, np.float64),
                  t(batch.scene_off, np.int32), t(self.init_h, np.float64)]
        self.d_out = torch.zeros(self.n_out * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        self.cells_hint = capi.max_cells(batch.model_off)     # what the host-buffer call computes itself from the offsets
        # inputs as the API delivers them (FP64 xy) read once + offsets + init + record
        self.alg_bytes = 16 * (len(batch.model_xy) + len(batch.scene_xy)) + self.n_pairs * 16 + self.n_pairs * n_seeds * 48 + self.n_out * 160
        self.h2d = int(batch.model_xy.nbytes + batch.scene_xy.nbytes + batch.model_off.nbytes + batch.scene_off.nbytes + self.init_h.nbytes)
        self.d2h = int(self.out_h.nbytes)

    def launch(self, ctx, stream, prm=None):
        d = self.d
        ctx.register_batch_device(self.n_pairs, self.n_seeds, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(),
                                  d[4].data_ptr(), prm or self.prm, self.d_out.data_ptr(), 0, self.batch.max_model,
                                  self.batch.max_scene, stream.cuda_stream, max_cells_hint=self.cells_hint)

    def host_call(self, ctx):
        b = self.batch
        ctx.register_batch(b.model_xy, b.model_off, b.scene_xy, b.scene_off, self.init_h, self.prm, n_seeds=self.n_seeds, out=self.out_h)


def time_legs(torch, ctx, db, stream, steps, warmup, barrier, gather=None, e2e=True):
    """Device-resident leg (CUDA events on the launching stream around `steps` steps) and, optionally, the end-to-end
    leg through the host-buffer call.  Returns (dev_ms, kernel_ms mean, launches per device step, e2e_ms, e2e launches)."""
    with torch.cuda.stream(stream):
        for _ in range(warmup):
            db.launch(ctx, stream)
            if gather:
                gather.step()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kernel_ms = []
        ev0.record(stream)
        for _ in range(steps):
            db.launch(ctx, stream)
            if gather:
                gather.step()             # gather of this step's records on a side stream, under the next step's kernel
            stream.synchronize()          # per-step so the library's own kernel events can be read back
            kernel_ms.append(ctx.launch_info().last_kernel_ms)
        if gather:
            gather.drain()                # the timed region ends when the last gather has landed
        ev1.record(stream)
        barrier()
        dev_ms = ev0.elapsed_time(ev1) / steps
        li = ctx.launch_info()
        e2e_ms, e2e_launches = None, 0
        if e2e:
            for _ in range(max(1, min(warmup, 2))):
                db.host_call(ctx)
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                db.host_call(ctx)
            torch.cuda.synchronize()
            e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
            e2e_launches = int(ctx.launch_info().n_launches)
            barrier()
    return dev_ms, float(np.mean(kernel_ms)), li, e2e_ms, e2e_launches


def sub_config(torch, dev, ctx, capi, name, workload, batch, init_Rt, n_seeds, prm, steps, warmup, kernel, unit_name, traffic,
               peak, peak_src, cpu):
    """One of the non-headline configs (C4, C5; C3 under torchrun has its own path) on this GPU, same legs as the headline."""
    db = DeviceBatch(torch, dev, capi, batch, init_Rt, n_seeds, prm)
    stream = torch.cuda.Stream(device=dev)
    barrier = torch.cuda.synchronize
    dev_ms, k_ms, li, e2e_ms, e2e_l = time_legs(torch, ctx, db, stream, steps, warmup, barrier)
    units = db.n_pairs * n_seeds
    kernel = {0: "icp_pair_seeds_kernel" if n_seeds > 1 else "icp_pair_kernel", 1: "icp_pair_l2_kernel",
              4: "icp_pair_cluster_kernel"}.get(int(li.kernel_variant), kernel)
    dram, _, cap = kernel_traffic(traffic, kernel, db.n_pairs)
    ach = db.alg_bytes / (k_ms * 1e-3) / 1e9
    rec = {"workload": workload, "value": units / (dev_ms * 1e-3), "unit": unit_name, "ms_per_step": dev_ms, "kernel_ms": k_ms,
           "steps": steps, "warmup": warmup, "kernel": kernel, "kernel_variant": int(li.kernel_variant), "grid": int(li.grid), "block": int(li.block),
           "launches_per_step": int(li.n_launches), "algorithmic_bytes": db.alg_bytes, "dram_bytes": dram, "dram_capture": cap,
           "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": dram,
                        "peak_source": peak_src},
           "e2e": {"value": units / (e2e_ms * 1e-3), "unit": unit_name, "ms_per_step": e2e_ms, "h2d_bytes_per_step": db.h2d,
                   "d2h_bytes_per_step": db.d2h, "launches_per_step": e2e_l},
           "cpu_baseline": cpu}
    rec["ok_frac"] = float(((db.d_out.cpu().numpy().view(capi.RESULT_DTYPE)["status"] & (capi.ST_TOO_FEW_MODEL | capi.ST_TOO_FEW_SCENE | capi.ST_HINT_TOO_SMALL)) == 0).mean())
    del db
    torch.cuda.empty_cache()
    return rec, int(li.n_launches) * steps + e2e_l * steps


def c5_inputs(n_pairs, n_seeds, pinned=None):
    from slam_b200 import scangen
    b = generate_chunked(C5_SEED0, n_pairs, 1, pinned) if pinned else scangen.generate(C5_SEED0, n_pairs, config=1)
    init = np.concatenate([scangen.pose_to_Rt(scangen.hypotheses(int(sd), n_seeds)) for sd in b.seeds], 0)
    return b, init


def gpu_arm(args, rank, world, local_rank, cpu_lines):
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
    prm = capi.fixed_iter_params(ITERS, indist=INDIST)
    ctx = capi.Context(local_rank)
    db = DeviceBatch(torch, dev, capi, batch, batch.init_Rt(), 1, prm)
    stream = torch.cuda.Stream(device=dev)
    gather = sharding.RecordGather(db.d_out, world, n_pairs * world, stream) if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # the path's only exchange: gather the per-pair records, overlapped with the next step's kernel
    dev_ms, k_ms, li, e2e_ms, e2e_launches = time_legs(torch, ctx, db, stream, args.steps, args.warmup, barrier,
                                                       gather=gather)
    if gather:
        gather.finish()
    clocks = sampler.stop() if rank == 0 else None

    # sanity: the device leg and the host leg computed the same records
    dev_rec = db.d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    out_h = db.out_h
    same = bool((dev_rec["num_corr"] == out_h["num_corr"]).all() and np.array_equal(dev_rec["R"], out_h["R"]))
    ok_mask = np.uint32(capi.ST_TIE_REPLAY | capi.ST_CONVERGED | capi.ST_EXACT_RETRY)
    ok_frac = float((out_h["status"] & ~ok_mask == 0).mean())

    # useful work of one launch (diagnostic instantiation, outside every timed region)
    useful = None
    if rank == 0:
        prm_d = capi.fixed_iter_params(ITERS, indist=INDIST, flags=capi.F_DIAG_COUNTS)
        ctx.diag_counts()
        db.launch(ctx, stream, prm=prm_d)
        stream.synchronize()
        useful = useful_roofline(ctx.diag_counts(), 1, k_ms)

    # the other configs of BASELINE.json, each through the same two legs
    configs, extra_launches = {}, 0
    traffic = load_traffic()
    peak, peak_src = load_peak()
    if world > 1 and not c3 and args.configs:
        # C3: the whole 262144-pair loop-closure batch split over the ranks (strong scaling), gather in the timed region
        del db
        torch.cuda.empty_cache()
        lo3, n3 = sharding.shard_seeds(C3_SEED0, args.total_pairs, rank, world)
        if n3 * world == args.total_pairs:
            b3 = generate_chunked(lo3, n3, 3, capi.pinned_array)
            db3 = DeviceBatch(torch, dev, capi, b3, b3.init_Rt(), 1, prm)
            g3 = sharding.RecordGather(db3.d_out, world, n3 * world, stream)
            d3, k3, li3, e3, el3 = time_legs(torch, ctx, db3, stream, args.sub_steps, 2, barrier, gather=g3)
            g3.finish()
            tm3 = torch.tensor([d3, e3], dtype=torch.float64, device=dev)
            dist.all_reduce(tm3, op=dist.ReduceOp.MAX)
            d3, e3 = tm3.tolist()
            dram3, _, cap3 = kernel_traffic(traffic, "icp_pair_kernel", n3)
            ach3 = db3.alg_bytes / (k3 * 1e-3) / 1e9
            configs["c3"] = {"workload": "C3: %d loop-closure scan pairs in total (%d per GPU), offsets +-2 m / +-0.3 rad, init = truth + N(0.3 m, "
                                         "0.05 rad), 1081 beams, %d fixed iterations; records all-gathered over NCCL inside the timed region"
                                         % (n3 * world, n3, ITERS),
                             "scaling": "strong", "value": n3 * world / (d3 * 1e-3), "unit": UNIT, "ms_per_step": d3, "kernel_ms": k3,
                             "steps": args.sub_steps, "warmup": 2, "kernel": "icp_pair_kernel", "algorithmic_bytes": db3.alg_bytes * world,
                             "dram_bytes": dram3, "dram_capture": cap3,
                             "roofline": {"bound": "hbm", "achieved": ach3, "peak": peak, "unit": "GB/s", "frac": ach3 / peak, "traffic": dram3,
                                          "peak_source": peak_src, "note": "per GPU"},
                             "e2e": {"value": n3 * world / (e3 * 1e-3), "unit": UNIT, "ms_per_step": e3, "h2d_bytes_per_step": db3.h2d * world,
                                     "d2h_bytes_per_step": db3.d2h * world},
                             "cpu_baseline": None, "cpu_baseline_note": "measured at N=1 only (rank 0), see the 1-GPU line"}
            extra_launches += (int(li3.n_launches) + el3) * args.sub_steps * world
            del db3, g3
    elif world == 1 and not c3 and args.configs:
        del db
        torch.cuda.empty_cache()
        b4 = generate_chunked(C4_SEED0, args.c4_pairs, 4, capi.pinned_array, chunk=64)
        configs["c4"], n4 = sub_config(
            torch, dev, ctx, capi, "c4", "C4: %d pairs of dense 2-D slices, 32768 beams per scan (%d..%d points per cloud), %d fixed iterations, indist %g"
            % (b4.n_pairs, int((b4.model_off[2::2] - b4.model_off[:-2:2]).min()), b4.max_model, ITERS, INDIST),
            b4, b4.init_Rt(), 1, prm, args.sub_steps, 2, None, UNIT, traffic, peak, peak_src, cpu_lines.get("c4"))
        del b4
        b5, init5 = c5_inputs(args.c5_pairs, C5_SEEDS, capi.pinned_array)
        prm5 = capi.fixed_iter_params(ITERS, indist=INDIST, flags=capi.F_BEST_OF_SEEDS)
        configs["c5"], n5 = sub_config(
            torch, dev, ctx, capi, "c5", "C5: %d pairs x %d initial pose seeds (seed 0 = identity, the rest uniform in +-10 m, 0..2 pi), %d fixed "
            "iterations each, best seed selected on device" % (b5.n_pairs, C5_SEEDS, ITERS),
            b5, init5, C5_SEEDS, prm5, args.sub_steps, 2, "icp_pair_seeds_kernel", "problems/s", traffic, peak, peak_src, cpu_lines.get("c5"))
        configs["c5"]["pairs_per_s"] = configs["c5"]["value"] / C5_SEEDS
        extra_launches += n4 + n5

    tm = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = tm.tolist()
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return None

    alg_bytes = 16 * (len(batch.model_xy) + len(batch.scene_xy)) + n_pairs * (16 + 48 + 160)
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    dram, _, cap = kernel_traffic(traffic, "icp_pair_kernel", n_pairs)
    h2d = int(batch.model_xy.nbytes + batch.scene_xy.nbytes + batch.model_off.nbytes + batch.scene_off.nbytes + n_pairs * 48)
    d2h = int(n_pairs * 160)
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
                   "collective": "nccl all_gather of 160 B records, double-buffered: the gather of step k runs on a side stream "
                                 "under the kernel of step k+1" if world > 1 else "none"},
        "e2e": {"value": world * n_pairs / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                "d2h_bytes_per_step": d2h * world, "ms_per_step": e2e_ms,
                "api": "ccicp_register_batch (C ABI, pinned host buffers)", "host_numa": numa},
        # kernels of this library launched inside the timed regions (device leg + host-buffer leg, all configs), all ranks
        "gpu_launches": (int(li.n_launches) + e2e_launches) * args.steps * world + extra_launches,
        "gpu_launches_detail": {"device_leg_per_step": int(li.n_launches), "e2e_leg_per_step": e2e_launches,
                                "other_configs_total": extra_launches,
                                "note": "per ICP launch: icp_pair_kernel + collect_retry_kernel + the exact-order re-run kernel (usually empty)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": dram, "kernel": "icp_pair_kernel", "kernel_ms": k_ms, "algorithmic_bytes_per_launch": alg_bytes,
                     "peak_source": peak_src, "traffic_stale": bool(traffic and traffic.get("stale")), "traffic_match": (traffic or {}).get("match"),
                     "note": "contract field.  The kernel keeps a pair on chip and moves only the algorithmic bytes, so HBM is not what "
                             "binds it (SURVEY.md 8(d)): roofline_useful is the fraction the >= 60 % target is about, roofline_issue the "
                             "scheduler occupancy"},
        # what binds this on-chip kernel: un-fused FP32 issue.  Useful operations only (counted by a diagnostic launch).
        "roofline_useful": useful,
        # scheduler occupancy: warp instructions of the committed ncu capture / live kernel time (stale-checked)
        "roofline_issue": issue_roofline(traffic, "icp_pair_kernel", n_pairs, k_ms, li.sm_count, clocks),
        "kernel": {"grid": li.grid, "block": li.block, "smem_bytes": li.smem_bytes, "regs": li.regs_per_thread,
                   "ctas_per_sm": li.blocks_per_sm, "sms": li.sm_count},
        "clocks": clocks,
        "check": {"device_leg_equals_host_leg": same, "pairs_status_ok_frac": ok_frac},
        "configs": configs,
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
    ap.add_argument("--no-configs", dest="configs", action="store_false",
                    help="skip the `configs` sub-records (C4 and C5 at 1 GPU, C3 under torchrun)")
    ap.add_argument("--sub-steps", type=int, default=3, help="timed steps of each `configs` sub-record")
    ap.add_argument("--c4-pairs", type=int, default=C4_PAIRS)
    ap.add_argument("--c5-pairs", type=int, default=C5_PAIRS)
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

    cpu_line, cpu_lines = None, {}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before any CUDA initialisation: the workers are forked
        from slam_b200 import scangen
        cores = host_cores()
        c3 = args.workload == "c3"
        sample = scangen.generate(C3_SEED0 if c3 else SEED0, min(args.pairs, 256), config=3 if c3 else 1)
        ppc = args.cpu_pairs_per_core or 1000        # ~12.5 ms each -> ~12 s of CPU work per core
        cpu_line, _ = cpu_reference_rate(sample, cores, ppc)
        if args.configs and not c3:
            s4 = scangen.generate(C4_SEED0, min(args.c4_pairs, cores), config=4)       # ~0.6 s per pair
            cpu_lines["c4"], _ = cpu_reference_rate(s4, cores, 2)
            s5, i5 = c5_inputs(min(args.c5_pairs, 8), C5_SEEDS)
            cpu_lines["c5"], _ = cpu_reference_rate(s5, cores, 128, inits=i5, seeds_per_pair=C5_SEEDS, unit="problems/s",
                                                    what="(pair, seed) problems")
    line = gpu_arm(args, rank, world, local_rank, cpu_lines)
    if line is not None:
        line["cpu_baseline"] = cpu_line
        line["bench_wall_s"] = round(time.perf_counter() - T_START, 1)
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
