"""ORACLE (test infrastructure, not product code) -- ggml block dequantisation.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package. The product path (llama_cpp/ + csrc/) never does.

PARITY UNPINNED BY THE REFERENCE: the arithmetic lives in llama-cpp-python==0.2.77
(pinned at /root/reference/docker/Dockerfile.base:30, /root/reference/README.md:20),
which vendors llama.cpp/ggml; that source is not under /root/reference, is not
installed here and cannot be fetched. The reference repo has no tests or golden
vectors. What this file IS pinned against: gguf-py 0.19
(`gguf.quants.dequantize`, ggml's own numpy reference, present in this image) --
tests/test_oracle_blocks.py checks bit-equality on random and hand-built blocks,
and tests/golden/blocks_*.npz holds committed vectors generated from it.

Published block layouts restated here (all little endian, fp16 scale fields):

  Q4_0  (32 elems, 18 B): d | qs[16]; elem j<16 = low nibble of qs[j], elem 16+j = high
        nibble of qs[j]; value = (q - 8) * d                  [gguf/quants.py:241-251]
  Q8_0  (32 elems, 34 B): d | int8 qs[32]; value = q * d      [gguf/quants.py:396-401]
  Q4_K  (256 elems, 144 B): d | dmin | scales[12] | qs[128]; 8 sub-blocks of 32, 6-bit
        (scale, min) pairs packed in `scales`; per 64 elems, 32 bytes: low nibbles then
        high nibbles; value = (d*sc)*q - (dmin*m)             [gguf/quants.py:475-522]
  Q5_K  (256 elems, 176 B): d | dmin | scales[12] | qh[32] | qs[128]; as Q4_K with a 5th
        bit: bit s of qh[l] for sub-block s                   [gguf/quants.py:525-549]
  Q6_K  (256 elems, 210 B): ql[128] | qh[64] | int8 scales[16] | d; per 128 elems: low
        nibbles of ql[0..63], high nibbles of ql[0..63], 2-bit planes of qh[0..31];
        value = (d*scale[e/16]) * (q - 32)                    [gguf/quants.py:552-572]

  Q4_1  (32 elems, 20 B): d | m | qs[16]; nibbles as Q4_0; value = q * d + m            [gguf/quants.py Q4_1]
  Q5_0  (32 elems, 22 B): d | qh[4] | qs[16]; 5th bit of elem j = bit j of qh (u32);
        value = ((nibble | bit << 4) - 16) * d                                          [gguf/quants.py Q5_0]
  Q5_1  (32 elems, 24 B): d | m | qh[4] | qs[16]; value = (nibble | bit << 4) * d + m   [gguf/quants.py Q5_1]
  Q2_K  (256 elems, 84 B): scales[16] (low nibble scale, high nibble min, per 16 elems) | qs[64] | d | dmin;
        per 128 elems: 2-bit planes of qs[0..31]; value = (d*sc)*q - (dmin*m)           [gguf/quants.py Q2_K]
  Q3_K  (256 elems, 110 B): hmask[32] | qs[64] | scales[12] (16 x 6 bit, minus 32) | d;
        q = 2-bit plane of qs - (hmask bit ? 0 : 4); value = (d*scale) * q              [gguf/quants.py Q3_K]

All arithmetic is float32 with separate multiply and subtract (no fused multiply-add),
which is what the CUDA dequantisation kernels reproduce bit for bit.
"""
from __future__ import annotations

import numpy as np

F32, F16, Q4_0, Q8_0, Q4_K, Q5_K, Q6_K = 0, 1, 2, 8, 12, 13, 14
Q4_1, Q5_0, Q5_1, Q2_K, Q3_K = 3, 6, 7, 10, 11
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q8_0: (32, 34),
         Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210),
         Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q2_K: (256, 84), Q3_K: (256, 110)}


def _f16(b: np.ndarray) -> np.ndarray:
    """(n, 2) uint8 -> (n, 1) float32"""
    return np.ascontiguousarray(b).view(np.float16).astype(np.float32)


def k_scale_min(scales: np.ndarray):
    """Unpack the 12-byte Q4_K/Q5_K scale field -> (sc[n,8], m[n,8]) uint8.

    bytes 0-3: low 6 bits = sc[0..3], high 2 bits = bits 4-5 of sc[4..7]
    bytes 4-7: low 6 bits = m[0..3],  high 2 bits = bits 4-5 of m[4..7]
    bytes 8-11: low nibble = bits 0-3 of sc[4..7], high nibble = bits 0-3 of m[4..7]
    """
    s = scales.astype(np.uint8)
    a, b, c = s[:, 0:4], s[:, 4:8], s[:, 8:12]
    sc = np.concatenate([a & 63, (c & 15) | ((a >> 6) << 4)], axis=1)
    m = np.concatenate([b & 63, (c >> 4) | ((b >> 6) << 4)], axis=1)
    return sc, m


def dequantize(raw: np.ndarray, ggml_type: int) -> np.ndarray:
    """raw uint8 bytes (whole number of blocks) -> flat float32 values."""
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    be, bb = BLOCK[ggml_type]
    if ggml_type == F32:
        return raw.view(np.float32).copy()
    if ggml_type == F16:
        return raw.view(np.float16).astype(np.float32)
    blk = raw.reshape(-1, bb)
    n = blk.shape[0]
    if ggml_type == Q4_0:
        d = _f16(blk[:, 0:2])
        qs = blk[:, 2:18]
        q = np.concatenate([qs & 15, qs >> 4], axis=1).astype(np.int8) - np.int8(8)
        return (d * q.astype(np.float32)).reshape(-1)
    if ggml_type == Q8_0:
        d = _f16(blk[:, 0:2])
        q = blk[:, 2:34].view(np.int8).astype(np.float32)
        return (q * d).reshape(-1)
    if ggml_type in (Q4_K, Q5_K):
        d = _f16(blk[:, 0:2])
        dmin = _f16(blk[:, 2:4])
        sc, m = k_scale_min(blk[:, 4:16])
        if ggml_type == Q4_K:
            qs = blk[:, 16:144].reshape(n, 4, 32)
            hi = None
        else:
            qh = blk[:, 16:48]                       # (n, 32)
            qs = blk[:, 48:176].reshape(n, 4, 32)
            hi = ((qh[:, None, :] >> np.arange(8, dtype=np.uint8)[None, :, None]) & 1)  # (n, 8, 32)
        q = np.stack([qs & 15, qs >> 4], axis=2).reshape(n, 8, 32)   # sub-block order
        if hi is not None:
            q = q | (hi << 4).astype(np.uint8)
        dl = (d * sc.astype(np.float32)).reshape(n, 8, 1)
        ml = (dmin * m.astype(np.float32)).reshape(n, 8, 1)
        return (dl * q.astype(np.float32) - ml).reshape(-1)
    if ggml_type == Q6_K:
        ql = blk[:, 0:128].reshape(n, 2, 64)
        qh = blk[:, 128:192].reshape(n, 2, 32)
        sc = blk[:, 192:208].view(np.int8).astype(np.float32)
        d = _f16(blk[:, 208:210])
        lo = np.stack([ql & 15, ql >> 4], axis=2).reshape(n, 2, 128)       # e' = nib*64 + byte
        h2 = np.stack([(qh >> s) & 3 for s in (0, 2, 4, 6)], axis=2).reshape(n, 2, 128)  # e' = k*32 + byte
        q = (lo | (h2 << 4)).astype(np.int8) - np.int8(32)
        dl = (d * sc).reshape(n, 16, 1)
        return (dl * q.reshape(n, 16, 16).astype(np.float32)).reshape(-1)
    if ggml_type == Q4_1:
        d, m = _f16(blk[:, 0:2]), _f16(blk[:, 2:4])
        qs = blk[:, 4:20]
        q = np.concatenate([qs & 15, qs >> 4], axis=1).astype(np.float32)
        return (d * q + m).reshape(-1)
    if ggml_type in (Q5_0, Q5_1):
        o = 2 if ggml_type == Q5_0 else 4
        d = _f16(blk[:, 0:2])
        qh = np.ascontiguousarray(blk[:, o:o + 4]).view(np.uint32)                       # (n, 1)
        bit = ((qh >> np.arange(32, dtype=np.uint32)[None, :]) & 1).astype(np.uint8)      # (n, 32)
        qs = blk[:, o + 4:o + 20]
        q = np.concatenate([qs & 15, qs >> 4], axis=1) | (bit << 4)
        if ggml_type == Q5_0:
            return (d * (q.astype(np.int8) - np.int8(16)).astype(np.float32)).reshape(-1)
        return (d * q.astype(np.float32) + _f16(blk[:, 2:4])).reshape(-1)
    if ggml_type == Q2_K:
        sc = blk[:, 0:16]
        qs = blk[:, 16:80].reshape(n, 2, 1, 32)
        d, dmin = _f16(blk[:, 80:82]), _f16(blk[:, 82:84])
        dl = (d * (sc & 15).astype(np.float32)).reshape(n, 16, 1)
        ml = (dmin * (sc >> 4).astype(np.float32)).reshape(n, 16, 1)
        q = ((qs >> np.array([0, 2, 4, 6], dtype=np.uint8).reshape(1, 1, 4, 1)) & 3).reshape(n, 16, 16).astype(np.float32)
        return (dl * q - ml).reshape(-1)
    if ggml_type == Q3_K:
        hmask = blk[:, 0:32].reshape(n, 1, 1, 32)
        qs = blk[:, 32:96].reshape(n, 2, 1, 32)
        scb = blk[:, 96:108]
        d = _f16(blk[:, 108:110])
        lsc = (scb[:, 0:8].reshape(n, 1, 8) >> np.array([0, 4], dtype=np.uint8).reshape(1, 2, 1)).reshape(n, 16) & 15
        hsc = (scb[:, 8:12].reshape(n, 1, 4) >> np.array([0, 2, 4, 6], dtype=np.uint8).reshape(1, 4, 1)).reshape(n, 16) & 3
        scales = ((lsc | (hsc << 4)).astype(np.int8) - np.int8(32)).astype(np.float32)
        dl = (d * scales).reshape(n, 16, 1)
        ql = ((qs >> np.array([0, 2, 4, 6], dtype=np.uint8).reshape(1, 1, 4, 1)) & 3).reshape(n, 16, 16)
        qh = ((hmask >> np.arange(8, dtype=np.uint8).reshape(1, 1, 8, 1)) & 1).reshape(n, 16, 16) ^ 1
        q = (ql.astype(np.int8) - (qh << 2).astype(np.int8)).astype(np.float32)
        return (dl * q).reshape(-1)
    raise ValueError(f"unsupported ggml type {ggml_type}")


def dequantize_tensor(raw: np.ndarray, ggml_type: int, shape) -> np.ndarray:
    """shape in ggml order (K, N) -> float32 array [N, K]; 1-D -> [K]."""
    flat = dequantize(raw, ggml_type)
    if len(shape) == 1:
        return flat.reshape(shape[0])
    return flat.reshape(shape[1], shape[0])
