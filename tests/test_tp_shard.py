"""Tensor-parallel sharding of quantised tensors (llama_cpp/tp_shard.py), on the CPU with the block oracle: the shards of
a column-parallel matrix stack to the full product, the partial products of a row-parallel matrix add up to it, and no
block is re-quantised (the shards are byte ranges of the original)."""
import numpy as np
import pytest

from util import TYPE_IDS


@pytest.mark.parametrize("tname", ["Q4_K", "Q6_K", "Q5_K", "Q8_0", "Q4_0"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_shards_reproduce_the_unsharded_matvec(tname, world):
    from llama_cpp import tp_shard
    from llama_cpp.synth import random_blocks
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    be, bb = gb.BLOCK[t]
    K, N = 256 * world * 2, 16 * world
    rng = np.random.default_rng(world + t)
    w = random_blocks(tname, N * K // be, 0.02, seed=7 + t).reshape(-1)
    x = rng.standard_normal(K)
    W = gb.dequantize(w, t).reshape(N, K).astype(np.float64)
    full = W @ x
    # column-parallel: rows split, outputs concatenate
    cols = []
    for r in range(world):
        d, shape = tp_shard.shard_tensor("blk.0.ffn_up.weight", w, t, (K, N), r, world)
        assert shape == (K, N // world) and d.size == w.size // world
        cols.append(gb.dequantize(np.ascontiguousarray(d), t).reshape(N // world, K).astype(np.float64) @ x)
    assert np.array_equal(np.concatenate(cols), full)
    # row-parallel: K split at block boundaries, partial sums add up
    part = np.zeros(N)
    seen = 0
    for r in range(world):
        d, shape = tp_shard.shard_tensor("blk.0.ffn_down.weight", w, t, (K, N), r, world)
        kl = K // world
        assert shape == (kl, N)
        seen += d.size
        # the shard is made of the original bytes: row n of the shard == bytes [r, r+1) * kl/be*bb of row n
        rb = K // be * bb
        assert np.array_equal(d.reshape(N, -1), w.reshape(N, rb)[:, r * (kl // be * bb):(r + 1) * (kl // be * bb)])
        part += gb.dequantize(np.ascontiguousarray(d), t).reshape(N, kl).astype(np.float64) @ x[r * kl:(r + 1) * kl]
    assert seen == w.size
    assert np.allclose(part, full, rtol=1e-12, atol=1e-12)
    # replicated tensors come back untouched
    d, shape = tp_shard.shard_tensor("blk.0.attn_norm.weight", w, t, (K, N), 1, world)
    assert d is w and shape == (K, N)


def test_uneven_splits_are_refused():
    from llama_cpp import tp_shard
    w = np.zeros(3 * 144, np.uint8)
    with pytest.raises(ValueError):
        tp_shard.shard_tensor("blk.0.ffn_down.weight", w, TYPE_IDS["Q4_K"], (256, 3), 0, 2)     # K = 256 cannot split in two at 256-element blocks
    with pytest.raises(ValueError):
        tp_shard.shard_tensor("blk.0.attn_q.weight", w, TYPE_IDS["Q4_K"], (256, 3), 0, 2)       # 3 rows over 2 ranks
