"""CPU check of the device block arithmetic: csrc/quant_blocks.cuh is compiled for the host
(B2_HOST_EMU) and compared with the oracle. Catches nibble / bit-plane / alignment mistakes
before any GPU time is spent."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import PKG
from util import P, TYPE_IDS, finite_random_blocks

QTYPES = ["Q4_0", "Q8_0", "Q4_K", "Q5_K", "Q6_K", "Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K"]


@pytest.fixture(scope="module")
def host_lib():
    out = os.path.join(PKG, "lib", "libb200_hosttest.so")
    src = os.path.join(PKG, "csrc", "host_selftest.cpp")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", src, "-o", out], check=True)
    return ctypes.CDLL(out)


@pytest.mark.parametrize("tname", QTYPES + ["F16", "F32"])
def test_host_dequant_bit_exact(host_lib, tname):
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    b = finite_random_blocks(tname, 500, np.random.default_rng(1))
    ref = gb.dequantize(b, t)
    out = np.empty_like(ref)
    assert host_lib.b2_host_dequant(t, b.ctypes.data_as(P), ctypes.c_long(ref.size), out.ctypes.data_as(P)) == 0
    assert np.array_equal(ref.view(np.uint32), out.view(np.uint32))


@pytest.mark.parametrize("tname", QTYPES)
@pytest.mark.parametrize("K", [256, 2048, 5632])
def test_host_chunk_dot_vs_oracle(host_lib, tname, K):
    from llama_cpp.synth import random_blocks
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    N = 8
    rng = np.random.default_rng(K)
    w = random_blocks(tname, N * K // gb.BLOCK[t][0], 0.02, seed=K + t).reshape(N, -1)
    x = (rng.standard_normal(K) * rng.uniform(0.1, 3, size=K)).astype(np.float32)
    W = gb.dequantize(w, t).reshape(N, K).astype(np.float64)
    ref = W @ x.astype(np.float64)
    for mis in {"Q6_K": (0, 2), "Q3_K": (0, 2, 14), "Q2_K": (0, 4, 12)}.get(tname, (0,)):
        y = np.empty(N, np.float32)
        assert host_lib.b2_host_matvec(t, w.ctypes.data_as(P), N, K, x.ctypes.data_as(P), y.ctypes.data_as(P), 1, mis) == 0
        assert np.abs(y - ref).max() / np.abs(ref).std() < 1e-3, (tname, K, mis)


def test_two_plane_activations_are_about_16_bits(host_lib):
    x = np.random.default_rng(0).standard_normal(4096).astype(np.float32)
    xr = np.empty_like(x)
    host_lib.b2_host_act_roundtrip(x.ctypes.data_as(P), 4096, xr.ctypes.data_as(P), 1)
    assert np.sqrt(((x - xr) ** 2).mean()) / x.std() < 1e-4
    host_lib.b2_host_act_roundtrip(x.ctypes.data_as(P), 4096, xr.ctypes.data_as(P), 0)
    assert np.sqrt(((x - xr) ** 2).mean()) / x.std() < 2e-2
