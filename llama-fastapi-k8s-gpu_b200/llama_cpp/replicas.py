"""Multi-GPU plumbing for models that fit one GPU: independent replicas, no data-path collective.

The reference scales out with 4 single-GPU pods behind a k8s Service
(/root/reference/helm/values.yaml:17, docker/Dockerfile.app:12); here a replica is one engine per
GPU (one process per GPU under torchrun). The only cross-rank traffic is bookkeeping: a barrier
around the timed region and a max-reduction of the per-rank times (bench.py), over NCCL on the GPU
box and over gloo in the CPU tests.
"""
from __future__ import annotations

from typing import List, Sequence


def shard_range(n_items: int, world: int, rank: int) -> range:
    """Contiguous, balanced split of n_items independent requests across ranks."""
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def dist_max(values: Sequence[float], device=None) -> List[float]:
    """Element-wise max over ranks (identity when torch.distributed is not initialised)."""
    try:
        import torch
        import torch.distributed as dist
    except Exception:
        return list(values)
    if not dist.is_available() or not dist.is_initialized():
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def dist_sum(values: Sequence[float], device=None) -> List[float]:
    try:
        import torch
        import torch.distributed as dist
    except Exception:
        return list(values)
    if not dist.is_available() or not dist.is_initialized():
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()]


def aggregate_throughput(tokens_this_rank: int, ms_this_rank: float) -> dict:
    """Whole-job tok/s = all ranks' tokens / the slowest rank's time (bench.py contract)."""
    ms = dist_max([ms_this_rank])[0]
    total = dist_sum([float(tokens_this_rank)])[0]
    return {"value": total / (ms / 1e3), "ms": ms, "tokens": int(total)}
