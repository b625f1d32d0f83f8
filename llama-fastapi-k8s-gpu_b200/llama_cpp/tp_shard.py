"""Tensor-parallel sharding of GGUF tensors (70B-class models, BASELINE.json configs[4]).

The quantised blocks are never re-quantised, only selected:

  * column-parallel -- attn_q / attn_k / attn_v, ffn_gate / ffn_up, output: the stored rows (ggml ne1) are split, rank r
    takes rows [r, r+1) * N / world. Rows are contiguous in the file, so a shard is one contiguous byte range. Heads are
    consecutive row ranges, so rank r ends up with query heads [r, r+1) * n_head / world and the KV heads that serve them.
  * row-parallel -- attn_output, ffn_down: the contraction dimension K (ggml ne0) is split at block boundaries
    (K / world must be a multiple of the type's block size, 256 for K-quants), rank r takes the same K range of EVERY
    row: a strided byte range, gathered into a contiguous buffer at upload. The partial products of the ranks add up to
    the full mat-vec (all-reduce in the engine).
  * everything else (norm vectors, token_embd) is replicated.

The reference has no tensor parallelism (replicas only, /root/reference/helm/values.yaml:17).
"""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np

from .gguf_file import GGML_TYPES

COLUMN_PARALLEL = ("attn_q.weight", "attn_k.weight", "attn_v.weight", "ffn_gate.weight", "ffn_up.weight")
ROW_PARALLEL = ("attn_output.weight", "ffn_down.weight")


def shard_kind(name: str) -> str:
    if name == "output.weight" or name.endswith(COLUMN_PARALLEL):
        return "column"
    if name.endswith(ROW_PARALLEL):
        return "row"
    return "replicated"


def shard_tensor(name: str, data: np.ndarray, ggml_type: int, shape: Tuple[int, ...], rank: int, world: int):
    """-> (uint8 array holding this rank's blocks, local ggml shape). `data` is the tensor's raw bytes (uint8)."""
    kind = shard_kind(name)
    if kind == "replicated" or world == 1:
        return data, tuple(shape)
    _, be, bb = GGML_TYPES[ggml_type]
    K, N = int(shape[0]), int(shape[1])
    row_bytes = K // be * bb
    rows = np.asarray(data, dtype=np.uint8).reshape(N, row_bytes)
    if kind == "column":
        if N % world:
            raise ValueError(f"{name}: {N} rows do not split over {world} ranks")
        n = N // world
        return rows[rank * n:(rank + 1) * n].reshape(-1), (K, n)
    if K % (world * be):
        raise ValueError(f"{name}: K = {K} does not split over {world} ranks at {be}-element block boundaries")
    kb = K // world // be * bb                       # bytes of one row's K range
    return np.ascontiguousarray(rows[:, rank * kb:(rank + 1) * kb]).reshape(-1), (K // world, N)


def local_dims(dims, world: int):
    """Head counts and n_ff of one rank (n_embd and n_vocab stay global)."""
    import dataclasses
    if dims.n_head % world or dims.n_head_kv % world or dims.n_ff % world:
        raise ValueError(f"n_head {dims.n_head}, n_head_kv {dims.n_head_kv}, n_ff {dims.n_ff} must all divide by tp world {world}")
    if (dims.n_ff // world) % 256 or (dims.n_head // world * dims.head_dim) % 256:
        raise ValueError("the per-rank n_ff and n_head * head_dim must stay multiples of 256")
    if dims.n_vocab % (2 * world):
        raise ValueError("n_vocab must be a multiple of 2 * tp world")
    return dataclasses.replace(dims, n_head=dims.n_head // world, n_head_kv=dims.n_head_kv // world, n_ff=dims.n_ff // world)


def shard_all(tensors: Dict[str, object], rank: int, world: int):
    """name -> (data, ggml_type, local shape) for every tensor a rank uploads; a tied output matrix gets its own shard."""
    out = {}
    for name, t in tensors.items():
        data, shape = shard_tensor(name, t.data, t.ggml_type, t.shape, rank, world)
        out[name] = (data, t.ggml_type, shape)
    if world > 1 and "output.weight" not in tensors:
        t = tensors["token_embd.weight"]
        data, shape = shard_tensor("output.weight", t.data, t.ggml_type, t.shape, rank, world)
        out["output.weight"] = (data, t.ggml_type, shape)
    return out
