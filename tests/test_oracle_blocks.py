"""The block oracle is pinned against ggml's own numpy reference (gguf-py) and the committed
golden vectors; the reference repo has no vectors of its own."""
import os

import numpy as np
import pytest

from util import TYPE_IDS, finite_random_blocks

QTYPES = ["Q4_0", "Q8_0", "Q4_K", "Q5_K", "Q6_K", "Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K"]
GOLD = os.path.join(os.path.dirname(__file__), "golden", "blocks.npz")


@pytest.mark.parametrize("tname", QTYPES)
def test_oracle_matches_golden_vectors(tname):
    from oracle import ggml_blocks as gb
    g = np.load(GOLD)
    out = gb.dequantize(g[f"{tname}_raw"], TYPE_IDS[tname])
    assert np.array_equal(out.view(np.uint32), g[f"{tname}_f32"].view(np.uint32))


@pytest.mark.parametrize("tname", QTYPES)
def test_oracle_matches_gguf_py(tname):
    quants = pytest.importorskip("gguf.quants")
    from gguf.constants import GGMLQuantizationType as T
    from oracle import ggml_blocks as gb
    rng = np.random.default_rng(99)
    b = finite_random_blocks(tname, 2000, rng)
    ref = quants.dequantize(b, getattr(T, tname)).reshape(-1)
    out = gb.dequantize(b, TYPE_IDS[tname])
    assert np.array_equal(out.view(np.uint32), ref.view(np.uint32))


def test_hand_built_blocks():
    """Known-answer blocks written by hand from the published layouts."""
    from oracle import ggml_blocks as gb
    one = np.array([1.0], np.float16).view(np.uint8)
    # Q4_0: d=1, byte j = j | ((15-j) << 4) -> elem j = j-8, elem 16+j = 7-j
    b = np.zeros(18, np.uint8)
    b[0:2] = one
    b[2:18] = [j | ((15 - j) << 4) for j in range(16)]
    v = gb.dequantize(b, gb.Q4_0)
    assert v.tolist() == [j - 8 for j in range(16)] + [7 - j for j in range(16)]
    # Q8_0: d=0.5
    b = np.zeros(34, np.uint8)
    b[0:2] = np.array([0.5], np.float16).view(np.uint8)
    b[2:34] = np.arange(-16, 16, dtype=np.int8).view(np.uint8)
    assert gb.dequantize(b, gb.Q8_0).tolist() == [0.5 * q for q in range(-16, 16)]
    # Q4_K: d=1, dmin=1, sub-block s has scale s+1 and min s; all nibbles = 3
    b = np.zeros(144, np.uint8)
    b[0:2] = one
    b[2:4] = one
    sc = [s + 1 for s in range(8)]
    mn = list(range(8))
    for i in range(4):
        b[4 + i] = sc[i] | ((sc[4 + i] >> 4) << 6)
        b[8 + i] = mn[i] | ((mn[4 + i] >> 4) << 6)
        b[12 + i] = (sc[4 + i] & 15) | ((mn[4 + i] & 15) << 4)
    b[16:144] = 0x33
    v = gb.dequantize(b, gb.Q4_K).reshape(8, 32)
    for s in range(8):
        assert np.all(v[s] == (s + 1) * 3 - s)
    # Q6_K: d=1, scale g = g-8, ql nibbles 5, qh bits 2 -> q = (5 | 2<<4) - 32 = 5
    b = np.zeros(210, np.uint8)
    b[0:128] = 0x55
    b[128:192] = 0xAA
    b[192:208] = np.arange(-8, 8, dtype=np.int8).view(np.uint8)
    b[208:210] = one
    v = gb.dequantize(b, gb.Q6_K).reshape(16, 16)
    for g in range(16):
        assert np.all(v[g] == (g - 8) * 5)
    # Q5_K: as Q4_K with the fifth bit set for odd sub-blocks only
    b = np.zeros(176, np.uint8)
    b[0:2] = one
    for i in range(4):
        b[4 + i] = 1
        b[12 + i] = 1
    b[16:48] = 0xAA        # bit s of qh set for odd s
    b[48:176] = 0x21       # low nibble 1, high nibble 2
    v = gb.dequantize(b, gb.Q5_K).reshape(8, 32)
    for s in range(8):
        lo = 1 if s % 2 == 0 else 2
        assert np.all(v[s] == lo + (16 if s % 2 else 0))


def test_empty_and_ragged_input():
    from oracle import ggml_blocks as gb
    assert gb.dequantize(np.zeros(0, np.uint8), gb.Q4_K).size == 0
    with pytest.raises(ValueError):
        gb.dequantize(np.zeros(100, np.uint8), gb.Q4_K)
