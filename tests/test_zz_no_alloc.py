"""b2_attention_decode_static: the decode-attention entry point that never allocates.

It is for hosts that cannot ask the driver for memory when they call the library -- a server
that has already given all of its HBM to weights and KV pages, or a token loop that must not
touch the allocator (an allocation stalls the device and can fail when memory is tight). Its
split-merge scratch is a fixed array inside the library, shared by every call.

The test drives it through the C ABI with caller-owned device buffers (torch is only used here,
on the test side, to own those buffers) and checks the result against the existing entry point
b2_op_attention_decode, twice in a row: the second call must not be disturbed by what the first
left in the shared scratch. Collected last (file name) so a -x run has already run everything else.
"""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import PKG, ROOT
from util import lib_check, ptr

pytestmark = pytest.mark.gpu


def _paged_pool(kv, blocks, total, n_kv, bt, hd):
    """token-major [T, n_kv, hd] f16 -> paged pool [total][n_kv][bt][hd] in the order `blocks` gives"""
    T = kv.shape[0]
    pool = np.zeros((total, n_kv, bt, hd), np.float16)
    for t in range(T):
        pool[blocks[t // bt], :, t % bt, :] = kv[t]
    return pool


@pytest.mark.parametrize("n_head,n_kv,hd,bt,n_splits", [(32, 8, 128, 64, 6), (8, 2, 64, 16, 3), (32, 8, 128, 64, 2),
                                                        (32, 8, 128, 64, 4), (16, 4, 64, 32, 8)])
def test_static_scratch_attention_matches_existing_entry_point(native_lib, n_head, n_kv, hd, bt, n_splits):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(n_head + hd)
    dev = torch.device("cuda", 0)
    for call, T in enumerate((333, 170)):                      # two calls in a row, different inputs
        q = rng.standard_normal(n_head * hd).astype(np.float32)
        k = (rng.standard_normal((T, n_kv, hd)) * 0.7).astype(np.float16)
        v = rng.standard_normal((T, n_kv, hd)).astype(np.float16)
        nblk = (T + bt - 1) // bt
        total = nblk + 2
        blocks = rng.permutation(total)[:nblk].astype(np.int32)
        want = np.zeros(n_head * hd, np.float32)
        lib_check(native_lib, native_lib.b2_op_attention_decode(
            0, ptr(q), ptr(k.view(np.uint16)), ptr(v.view(np.uint16)), T, n_head, n_kv, hd, bt, ptr(blocks), total,
            n_splits, ptr(want)))
        # caller-owned device buffers
        d_q = torch.from_numpy(q).to(dev)
        d_k = torch.from_numpy(_paged_pool(k, blocks, total, n_kv, bt, hd)).to(dev)
        d_v = torch.from_numpy(_paged_pool(v, blocks, total, n_kv, bt, hd)).to(dev)
        d_tab = torch.from_numpy(blocks).to(dev)
        d_pos = torch.tensor([T - 1], dtype=torch.int32, device=dev)
        d_out = torch.full((n_head * hd,), float("nan"), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        rc = native_lib.b2_attention_decode_static(
            0, d_q.data_ptr(), d_k.data_ptr(), d_v.data_ptr(), d_tab.data_ptr(), nblk, d_pos.data_ptr(), n_head, n_kv, hd,
            bt, n_splits, d_out.data_ptr(), None)
        lib_check(native_lib, rc)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        # same tolerance as the attention parity tests (float32 accumulation over f16 K/V)
        assert np.abs(got - want).max() < 2e-4, (call, T, np.abs(got - want).max())


def test_static_scratch_limits_are_enforced(native_lib):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    buf = torch.zeros(1 << 16, dtype=torch.float32, device=dev)
    ints = torch.zeros(16, dtype=torch.int32, device=dev)
    p = buf.data_ptr()
    # 64 heads x 60 splits x 130 floats is over the 262144-float scratch: refused, with the limit in the message
    rc = native_lib.b2_attention_decode_static(0, p, p, p, ints.data_ptr(), 4, ints.data_ptr(), 64, 8, 128, 16, 60, p, None)
    assert rc != 0 and b"scratch" in native_lib.b2_last_error()
    # a single split needs no scratch and is not what this entry point is for
    assert native_lib.b2_attention_decode_static(0, p, p, p, ints.data_ptr(), 4, ints.data_ptr(), 8, 2, 64, 16, 1, p, None) != 0
    assert native_lib.b2_attention_decode_static(0, None, p, p, ints.data_ptr(), 4, ints.data_ptr(), 8, 2, 64, 16, 4, p, None) != 0


_COLD = """
import ctypes, sys
import numpy as np
import torch
sys.path[:0] = [{root!r}, {pkg!r}, {tests!r}]
from llama_cpp import _cabi
from util import ptr
lib = _cabi.load_library()
n_head, n_kv, hd, bt, n_splits = 32, 8, 128, 64, 6
rng = np.random.default_rng(5)
dev = torch.device("cuda", 0)
free = []
outs = []
cases = []
for T in (200, 421):
    q = rng.standard_normal(n_head * hd).astype(np.float32)
    k = (rng.standard_normal((T, n_kv, hd)) * 0.7).astype(np.float16)
    v = rng.standard_normal((T, n_kv, hd)).astype(np.float16)
    nblk = (T + bt - 1) // bt
    blocks = rng.permutation(nblk + 2)[:nblk].astype(np.int32)
    def pool(x):
        p = np.zeros((nblk + 2, n_kv, bt, hd), np.float16)
        for t in range(T):
            p[blocks[t // bt], :, t % bt, :] = x[t]
        return torch.from_numpy(p).to(dev)
    bufs = (torch.from_numpy(q).to(dev), pool(k), pool(v), torch.from_numpy(blocks).to(dev),
            torch.tensor([T - 1], dtype=torch.int32, device=dev),
            torch.full((n_head * hd,), float("nan"), dtype=torch.float32, device=dev))
    cases.append((T, q, k, v, blocks, nblk, bufs))
torch.cuda.synchronize()
# the static entry point is the FIRST call into the library in this process
for T, q, k, v, blocks, nblk, (dq, dk, dv, dt, dp, do) in cases:
    before = torch.cuda.mem_get_info()[0]
    rc = lib.b2_attention_decode_static(0, dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dt.data_ptr(), nblk, dp.data_ptr(),
                                        n_head, n_kv, hd, bt, n_splits, do.data_ptr(), None)
    assert rc == 0, lib.b2_last_error()
    torch.cuda.synchronize()
    free.append((before, torch.cuda.mem_get_info()[0]))
    outs.append(do.cpu().numpy())
# the second call (everything lazily loaded by then) must leave the free-memory figure untouched
assert free[1][0] == free[1][1], free
for (T, q, k, v, blocks, nblk, _), got in zip(cases, outs):
    want = np.zeros(n_head * hd, np.float32)
    rc = lib.b2_op_attention_decode(0, ptr(q), ptr(k.view(np.uint16)), ptr(v.view(np.uint16)), T, n_head, n_kv, hd, bt,
                                    ptr(blocks), nblk + 2, n_splits, ptr(want))
    assert rc == 0, lib.b2_last_error()
    err = float(np.abs(got - want).max())
    assert err < 2e-4, (T, err)
print("cold-ok")
"""


def test_static_entry_point_as_first_library_call_in_a_fresh_process(native_lib):
    """A host of the kind this entry point is for never creates an engine: its first call into the library is
    this one. Run exactly that in a new process (nothing warmed up by other entry points), twice, and check that
    the device's free memory does not move across the second call and that both results match the existing
    entry point (called afterwards)."""
    pytest.importorskip("torch")
    code = _COLD.format(root=ROOT, pkg=PKG, tests=os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", textwrap.dedent(code)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "cold-ok" in r.stdout, r.stdout + r.stderr


def test_static_entry_point_rejects_misaligned_pools(native_lib):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    buf = torch.zeros(1 << 16, dtype=torch.float32, device=dev)
    ints = torch.zeros(16, dtype=torch.int32, device=dev)
    p = buf.data_ptr()
    rc = native_lib.b2_attention_decode_static(0, p, p + 2, p, ints.data_ptr(), 4, ints.data_ptr(), 8, 2, 64, 16, 4, p, None)
    assert rc != 0 and b"aligned" in native_lib.b2_last_error()
