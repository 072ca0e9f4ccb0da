#!/usr/bin/env python
"""What the box's host side can deliver: N ranks (one per GPU) doing nothing but pinned H2D copies of the bench's
per-step input size (142 MB) at the same time.  The end-to-end leg of bench.py at N GPUs cannot beat this ceiling.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 profiles/h2d_ceiling.py
prints one JSON line (rank 0): GB/s per rank (min / mean / max) and aggregate."""
import json
import os

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nbytes = 142 * 1000 * 1000
h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
h.fill_(1)
d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
st = torch.cuda.Stream(device=dev)
reps = 30
with torch.cuda.stream(st):
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    st.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record(st)
    st.synchronize()
gbs = nbytes * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
t = torch.tensor([gbs], dtype=torch.float64, device=dev)
if world > 1:
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    vals = [float(o.item()) for o in out]
    dist.destroy_process_group()
else:
    vals = [gbs]
if rank == 0:
    print(json.dumps({"n_gpus": world, "bytes_per_copy": nbytes, "copies": reps, "gbs_per_rank_min": min(vals),
                      "gbs_per_rank_mean": sum(vals) / len(vals), "gbs_per_rank_max": max(vals), "gbs_aggregate": sum(vals),
                      "host_cpus": len(os.sched_getaffinity(0)),
                      "how": "pinned torch tensor -> device, 30 back-to-back cudaMemcpyAsync per rank on its own stream, CUDA events, all ranks concurrently"}))
