"""The N>1 path on CPU: world_size-2 gloo run of the host-side batch split and the
max-over-ranks / whole-job aggregation used by bench.py (no GPU involved; the per-shard
"work" here is the CPU checker so that shard results can be compared with the whole batch)."""
import os
import socket

import numpy as np
import pytest

import brpc_b200  # noqa: F401  (registers the package alias)
from brpc_b200 import sharding


def test_split_by_bytes_partitions_and_balances():
    rng = np.random.default_rng(0)
    ln = np.clip(np.round(np.exp(rng.normal(np.log(16384), 2.0, 20000))), 64, 4 << 20).astype(np.uint32)
    for parts in (1, 2, 4, 8):
        b = sharding.split_by_bytes(ln, parts)
        assert b[0] == 0 and b[-1] == len(ln) and (np.diff(b) >= 0).all()
        w = np.array([ln[b[r]:b[r + 1]].astype(np.uint64).sum() for r in range(parts)], dtype=np.float64)
        assert w.max() <= w.sum() / parts + float(ln.max()) + 16 * len(ln)
    assert list(sharding.split_by_bytes(np.zeros(0, np.uint32), 4)) == [0, 0, 0, 0, 0]
    assert list(sharding.split_by_bytes(np.array([5], np.uint32), 2))[-1] == 1


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    ln = rng.integers(0, 9000, 3000).astype(np.uint32)
    arena, off, ln = O.corpus_batch("words", ln)
    lo, hi = sharding.shard_for_rank(ln, rank, world)
    crc, _ = O.port().batch_crc32c(arena, off[lo:hi], ln[lo:hi], threads=1)
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, hi, crc))
    t = sharding.max_over_ranks(10.0 * (rank + 1))
    rate = sharding.aggregate_throughput(float(ln[lo:hi].sum()), 1.0 + rank)
    if rank == 0:
        whole, _ = O.port().batch_crc32c(arena, off, ln, threads=1)
        cat = np.concatenate([g[2] for g in gathered])
        ok = (cat == whole).all() and gathered[0][1] == gathered[1][0] and gathered[-1][1] == len(ln)
        q.put((bool(ok), t, rate, float(ln.sum())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_split_and_reduce():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, t, rate, total = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert t == 20.0                      # max over ranks
    assert rate == pytest.approx(total / 2.0)  # all units / slowest rank's time
