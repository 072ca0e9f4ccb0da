"""Multi-GPU plumbing of the path: scan pairs are independent, so a batch is split into
contiguous pair ranges, one per rank (one process per GPU), and the only exchange is one all-gather of the
fixed-size result records.  torch.distributed is used for exactly that (NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np

from . import capi


def shard_range(n_items, rank, world):
    """Contiguous, balanced split: rank r owns [lo, hi); sizes differ by at most one, larger shards first."""
    base, rem = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_seeds(seed0, n_pairs_total, rank, world):
    """Seeds of the pairs a rank generates/owns (pair p of the job has seed seed0 + p)."""
    lo, hi = shard_range(n_pairs_total, rank, world)
    return seed0 + lo, hi - lo


def gather_records(local, world, n_total=None):
    """All-gathers per-pair result records.  `local` is a uint8 torch tensor holding this rank's records
    (len = n_local * 160).  Equal shards use one all_gather_into_tensor; ragged shards are padded to the
    largest shard and trimmed.  Returns a uint8 tensor with the records of all ranks in pair order."""
    import torch
    import torch.distributed as dist
    rec = capi.RESULT_DTYPE.itemsize
    if world == 1:
        return local
    n_local = local.numel() // rec
    if n_total is None:
        n_total = n_local * world
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    n_max = max(hi - lo for lo, hi in sizes)
    if n_max * world == n_total:
        out = torch.empty(n_total * rec, dtype=torch.uint8, device=local.device)
        dist.all_gather_into_tensor(out, local)
        return out
    padded = torch.zeros(n_max * rec, dtype=torch.uint8, device=local.device)
    padded[:n_local * rec] = local
    buf = torch.empty(world * n_max * rec, dtype=torch.uint8, device=local.device)
    dist.all_gather_into_tensor(buf, padded)
    parts = [buf[r * n_max * rec: r * n_max * rec + (hi - lo) * rec] for r, (lo, hi) in enumerate(sizes)]
    return torch.cat(parts)


class RecordGather:
    """The per-step gather of bench.py, taken off the critical path: after the kernel of step k has been enqueued on the
    compute stream, step() snapshots the rank's records into one of two staging buffers (so the kernel of step k+1 may
    overwrite them) and all-gathers that snapshot on a side stream, i.e. under the kernel of step k+1.  drain() makes
    the compute stream wait for every gather issued so far (call it before the closing event of a timed region: the
    gathers belong to the region).  On CPU tensors (gloo tests) the same calls run synchronously."""

    def __init__(self, local, world, n_total, compute_stream=None):
        import torch
        self.local, self.world, self.n_total, self.cs = local, world, n_total, compute_stream
        self.cuda = local.is_cuda
        self.k = 0
        self.stage = [torch.empty_like(local) for _ in range(2)]
        self.out = [None, None]
        if self.cuda:
            self.side = torch.cuda.Stream(device=local.device)
            self.copied = [torch.cuda.Event() for _ in range(2)]
            self.gathered = [None, None]

    def step(self):
        import torch
        k = self.k
        self.k ^= 1
        if not self.cuda:
            self.stage[k].copy_(self.local)
            self.out[k] = gather_records(self.stage[k], self.world, self.n_total)
            return
        cs = self.cs or torch.cuda.current_stream()
        if self.gathered[k] is not None:
            cs.wait_event(self.gathered[k])            # the gather that last read this staging buffer
        with torch.cuda.stream(cs):
            self.stage[k].copy_(self.local, non_blocking=True)
        self.copied[k].record(cs)
        self.side.wait_event(self.copied[k])
        with torch.cuda.stream(self.side):
            self.out[k] = gather_records(self.stage[k], self.world, self.n_total)
            ev = torch.cuda.Event()
            ev.record(self.side)
        self.gathered[k] = ev

    def drain(self):
        import torch
        if self.cuda:
            (self.cs or torch.cuda.current_stream()).wait_stream(self.side)

    def finish(self):
        """Waits for everything and returns the records of the last step (all ranks, pair order)."""
        import torch
        if self.cuda:
            self.side.synchronize()
        return self.out[self.k ^ 1]


def records_view(t):
    """uint8 torch tensor (CPU) -> numpy structured array of ccicp_result records."""
    return t.cpu().numpy().view(capi.RESULT_DTYPE)
