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
import ctypes

import numpy as np
import pytest

from util import lib_check, ptr

pytestmark = pytest.mark.gpu


def _paged_pool(kv, blocks, total, n_kv, bt, hd):
    """token-major [T, n_kv, hd] f16 -> paged pool [total][n_kv][bt][hd] in the order `blocks` gives"""
    T = kv.shape[0]
    pool = np.zeros((total, n_kv, bt, hd), np.float16)
    for t in range(T):
        pool[blocks[t // bt], :, t % bt, :] = kv[t]
    return pool


@pytest.mark.parametrize("n_head,n_kv,hd,bt,n_splits", [(32, 8, 128, 64, 6), (8, 2, 64, 16, 3)])
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
