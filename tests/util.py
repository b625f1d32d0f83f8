"""Shared helpers for the tests (ctypes plumbing around the C ABI)."""
import ctypes

import numpy as np

P = ctypes.c_void_p
TYPE_IDS = {"F32": 0, "F16": 1, "Q4_0": 2, "Q4_1": 3, "Q5_0": 6, "Q5_1": 7, "Q8_0": 8, "Q2_K": 10, "Q3_K": 11, "Q4_K": 12, "Q5_K": 13, "Q6_K": 14}


def ptr(a):
    return None if a is None else a.ctypes.data_as(P)


def finite_random_blocks(tname, n_blocks, rng):
    """Fully random block bytes with the fp16 scale fields forced finite (no inf/nan)."""
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    be, bb = gb.BLOCK[t]
    b = rng.integers(0, 256, size=(n_blocks, bb), dtype=np.uint8)
    cols = {"Q4_0": [0], "Q8_0": [0], "Q4_K": [0, 2], "Q5_K": [0, 2], "Q6_K": [208], "F16": [0],
            "Q4_1": [0, 2], "Q5_0": [0], "Q5_1": [0, 2], "Q2_K": [80, 82], "Q3_K": [108]}.get(tname, [])
    for c in cols:
        h = b[:, c:c + 2].copy().view(np.uint16)
        h[(h & 0x7C00) == 0x7C00] &= 0xBFFF
        b[:, c:c + 2] = h.view(np.uint8)
    if tname == "F32":
        f = b.view(np.uint32)
        f[(f & 0x7F800000) == 0x7F800000] &= 0xBFFFFFFF
    return b


def lib_check(lib, rc):
    if rc != 0:
        raise RuntimeError(f"rc={rc}: {lib.b2_last_error().decode()}")
