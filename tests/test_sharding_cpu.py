"""Host-side multi-rank logic on CPU: world_size-2 and -3 gloo process groups exercise the pair sharding and the
record gather that bench.py runs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from slam_b200 import capi, sharding


def test_shard_range_is_a_partition():
    for n in (0, 1, 7, 4096, 262144, 1000003):
        for world in (1, 2, 3, 4, 8):
            edges = [sharding.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
    assert sharding.shard_seeds(10**6, 262144, 3, 8) == (10**6 + 3 * 32768, 32768)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n_total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(n_total, rank, world)
    rec = np.zeros(hi - lo, capi.RESULT_DTYPE)
    rec["num_corr"] = np.arange(lo, hi)            # stands in for "the record of pair p"
    rec["t"][:, 0] = np.arange(lo, hi) * 0.5
    rec["best_seed"] = rank
    local = torch.from_numpy(rec.view(np.uint8).copy())
    out = sharding.records_view(sharding.gather_records(local, world, n_total))
    ok = (len(out) == n_total and (out["num_corr"] == np.arange(n_total)).all()
          and (out["t"][:, 0] == np.arange(n_total) * 0.5).all())
    # bench.py's per-step gather object (double-buffered; synchronous on CPU tensors): three steps, the local records
    # change between steps like a kernel overwriting them, the result of the LAST step is what finish() returns
    g = sharding.RecordGather(local, world, n_total)
    for step in range(3):
        rec["iters"] = step
        local.copy_(torch.from_numpy(rec.view(np.uint8).copy()))
        g.step()
    g.drain()
    last = sharding.records_view(g.finish())
    ok = ok and len(last) == n_total and (last["iters"] == 2).all() and (last["num_corr"] == np.arange(n_total)).all()
    tm = torch.tensor([float(rank + 1)])
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)        # bench.py's max-over-ranks timing reduction
    q.put((rank, bool(ok), float(tm.item())))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(2, 64), (2, 33), (3, 100)])
def test_gather_records_gloo(world, n_total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == list(range(world))
    assert all(r[1] for r in res) and all(r[2] == float(world) for r in res)


def test_bench_c3_shards_and_chunked_generation():
    """bench.py --workload c3: the ranks' shards of the loop-closure batch are exactly the batch (same seeds, same
    scans, same noisy initial poses), and the chunked generator packs the same CSR arrays as one call."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from slam_b200 import scangen
    plain = lambda shape, dt: np.zeros(shape, dt)
    total, world = 24, 3
    whole = scangen.generate(bench.C3_SEED0, total, config=3)
    parts = []
    for r in range(world):
        lo, n = sharding.shard_seeds(bench.C3_SEED0, total, r, world)
        parts.append(bench.generate_chunked(lo, n, 3, plain, chunk=3))       # 8 pairs per rank in chunks of 3, 3, 2
    assert np.array_equal(np.concatenate([p.seeds for p in parts]), whole.seeds)
    assert np.array_equal(np.concatenate([p.model_xy for p in parts]), whole.model_xy)
    assert np.array_equal(np.concatenate([p.scene_xy for p in parts]), whole.scene_xy)
    assert np.array_equal(np.concatenate([p.init for p in parts]), whole.init)
    off = np.concatenate([[0]] + [p.model_off[1:] + sum(len(q.model_xy) for q in parts[:i]) for i, p in enumerate(parts)])
    assert np.array_equal(off, whole.model_off)
    # C3's initial poses are the truth plus noise, not the identity of C2
    assert np.abs(whole.init - whole.truth).max() > 0 and np.abs(whole.init).max() > 0


def test_c_abi_shard_rule_matches_python():
    """ccicp_shard_range (C ABI, multi-device entry points) splits exactly like sharding.shard_range."""
    import ctypes as C
    lib = capi.load()
    lib.ccicp_shard_range.argtypes = [C.c_int32, C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.ccicp_shard_range.restype = None
    for n in (0, 1, 7, 97, 4096, 262144, 1000003):
        for world in (1, 2, 3, 4, 8):
            for r in range(world):
                b, e = C.c_int32(), C.c_int32()
                lib.ccicp_shard_range(n, world, r, C.byref(b), C.byref(e))
                assert (b.value, e.value) == sharding.shard_range(n, r, world)
