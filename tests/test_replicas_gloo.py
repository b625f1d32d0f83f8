"""world_size-2 gloo run of the replica plumbing bench.py uses at N > 1 (CPU only)."""
import os
import socket

import pytest


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from llama_cpp.replicas import aggregate_throughput, dist_max, shard_range
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard_range(11, world, rank)
        dist.barrier()
        agg = aggregate_throughput(tokens_this_rank=100 * (rank + 1), ms_this_rank=10.0 * (rank + 1))
        mx = dist_max([float(rank), 5.0 - rank])
        out.put((rank, list(mine), agg, mx))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_aggregation():
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    (r0, s0, a0, m0), (r1, s1, a1, m1) = res
    assert s0 == list(range(0, 6)) and s1 == list(range(6, 11))           # disjoint, covering, balanced
    assert a0 == a1 and a0["tokens"] == 300 and a0["ms"] == 20.0          # sum of tokens / slowest rank
    assert abs(a0["value"] - 300 / 0.020) < 1e-6
    assert m0 == m1 == [1.0, 5.0]


def test_single_process_is_identity():
    from llama_cpp.replicas import aggregate_throughput, shard_range
    assert list(shard_range(5, 1, 0)) == [0, 1, 2, 3, 4]
    assert aggregate_throughput(10, 5.0) == {"value": 2000.0, "ms": 5.0, "tokens": 10}
