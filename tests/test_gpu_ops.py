"""Parity of each CUDA kernel against the CPU oracle, called through the C ABI (GPU only).

Bars: block dequantisation bit-exact; mat-vec / attention within the
floating-point tolerances written next to each check; sampler: identical token.
"""
import ctypes

import numpy as np
import pytest

from util import P, TYPE_IDS, finite_random_blocks, lib_check, ptr

pytestmark = pytest.mark.gpu
GEMM_QTYPES = ["Q4_0", "Q8_0", "Q4_K", "Q5_K", "Q6_K"]        # block formats with a hand-written expansion in the GEMM
QTYPES = GEMM_QTYPES + ["Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K"]  # + the ones the GEMM dequantises element by element


@pytest.mark.parametrize("tname", QTYPES + ["F16", "F32"])
def test_dequantize_bit_exact(native_lib, tname):
    from oracle import ggml_blocks as gb
    rng = np.random.default_rng(7)
    t = TYPE_IDS[tname]
    be, bb = gb.BLOCK[t]
    for n_blocks in (1, 3, 1000, 4099):
        b = finite_random_blocks(tname, n_blocks, rng)
        ref = gb.dequantize(b, t)
        out = np.empty(n_blocks * be, np.float32)
        lib_check(native_lib, native_lib.b2_op_dequantize(0, t, ptr(b), n_blocks * be, ptr(out)))
        assert np.array_equal(ref.view(np.uint32), out.view(np.uint32)), (tname, n_blocks)
    # empty input is accepted
    lib_check(native_lib, native_lib.b2_op_dequantize(0, t, ptr(b), 0, ptr(out)))
    # ragged input (not a whole block) is rejected
    if be > 1:
        assert native_lib.b2_op_dequantize(0, t, ptr(b), be + 1, ptr(out)) != 0


def test_dequantize_golden(native_lib):
    import os
    from oracle import ggml_blocks as gb
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "blocks.npz"))
    for tname in QTYPES:
        raw, ref = g[f"{tname}_raw"], g[f"{tname}_f32"]
        out = np.empty_like(ref)
        lib_check(native_lib, native_lib.b2_op_dequantize(0, TYPE_IDS[tname], ptr(raw), ref.size, ptr(out)))
        assert np.array_equal(ref.view(np.uint32), out.view(np.uint32)), tname


def _matvec_ref(w, t, N, K, x, nw, eps, resid):
    from oracle import ggml_blocks as gb
    W = gb.dequantize(w, t).reshape(N, K).astype(np.float64)
    xx = x.astype(np.float64)
    if nw is not None:
        xx = xx / np.sqrt((xx ** 2).mean() + eps) * nw.astype(np.float64)
    y = W @ xx
    if resid is not None:
        y = y + resid.astype(np.float64)
    return y, np.sqrt(((W @ xx) ** 2).mean())


@pytest.mark.parametrize("tname", QTYPES)
@pytest.mark.parametrize("K,N", [(512, 2), (512, 150), (2048, 64), (4096, 600), (5632, 40), (14336, 300)])
def test_matvec_vs_oracle(native_lib, tname, K, N):
    from llama_cpp.synth import random_blocks
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    be = gb.BLOCK[t][0]
    rng = np.random.default_rng(K + N)
    w = random_blocks(tname, N * K // be, 0.02, seed=K * 31 + N).reshape(-1)
    x = (rng.standard_normal(K) * rng.uniform(0.2, 2.0, size=K)).astype(np.float32)
    nw = rng.uniform(0.5, 1.5, size=K).astype(np.float32)
    resid = rng.standard_normal(N).astype(np.float32)
    for use_norm, use_res, n_ctas in ((False, False, 0), (True, False, 0), (True, True, 296), (False, True, 5)):
        y = np.zeros(N, np.float32)
        rc = native_lib.b2_op_matvec(0, t, ptr(w), N, K, ptr(x), ptr(nw) if use_norm else None, 1e-5,
                                     ptr(resid) if use_res else None, ptr(y), 2, n_ctas)
        lib_check(native_lib, rc)
        ref, rms = _matvec_ref(w, t, N, K, x, nw if use_norm else None, 1e-5, resid if use_res else None)
        err = np.abs(y - ref).max() / rms
        # two int8 activation planes (~16 bits) + float32 accumulation: 1e-3 of the output rms
        assert err < 1e-3, (tname, K, N, use_norm, use_res, n_ctas, err)


def test_matvec_single_plane_is_looser_but_close(native_lib):
    from llama_cpp.synth import random_blocks
    K, N = 4096, 64
    rng = np.random.default_rng(3)
    w = random_blocks("Q4_K", N * K // 256, 0.02, seed=5).reshape(-1)
    x = rng.standard_normal(K).astype(np.float32)
    y = np.zeros(N, np.float32)
    lib_check(native_lib, native_lib.b2_op_matvec(0, 12, ptr(w), N, K, ptr(x), None, 0.0, None, ptr(y), 1, 0))
    ref, rms = _matvec_ref(w, 12, N, K, x, None, 0.0, None)
    assert np.abs(y - ref).max() / rms < 5e-2


def test_matvec_rejects_bad_shapes(native_lib):
    x = np.zeros(512, np.float32)
    y = np.zeros(4, np.float32)
    w = np.zeros(4 * 512, np.uint8)
    assert native_lib.b2_op_matvec(0, 12, ptr(w), 3, 512, ptr(x), None, 0.0, None, ptr(y), 2, 0) != 0   # odd rows
    assert native_lib.b2_op_matvec(0, 12, ptr(w), 4, 500, ptr(x), None, 0.0, None, ptr(y), 2, 0) != 0   # K % 256
    assert native_lib.b2_op_matvec(0, 99, ptr(w), 4, 512, ptr(x), None, 0.0, None, ptr(y), 2, 0) != 0   # bad type
    assert b"" != native_lib.b2_last_error()


@pytest.mark.parametrize("tg,tu", [("Q4_K", "Q4_K"), ("Q4_0", "Q4_0"), ("Q8_0", "Q8_0"), ("Q4_K", "Q6_K")])
@pytest.mark.parametrize("K,N", [(512, 1536), (4096, 1000), (2048, 5632)])
def test_gate_up_vs_oracle(native_lib, tg, tu, K, N):
    from llama_cpp.synth import random_blocks
    from oracle import ggml_blocks as gb
    rng = np.random.default_rng(11)
    wg = random_blocks(tg, N * K // gb.BLOCK[TYPE_IDS[tg]][0], 0.02, seed=1).reshape(-1)
    wu = random_blocks(tu, N * K // gb.BLOCK[TYPE_IDS[tu]][0], 0.02, seed=2).reshape(-1)
    x = rng.standard_normal(K).astype(np.float32)
    nw = rng.uniform(0.5, 1.5, size=K).astype(np.float32)
    y = np.zeros(N, np.float32)
    lib_check(native_lib, native_lib.b2_op_gate_up(0, TYPE_IDS[tg], ptr(wg), TYPE_IDS[tu], ptr(wu), N, K, ptr(x),
                                                   ptr(nw), 1e-5, ptr(y), 0))
    g, _ = _matvec_ref(wg, TYPE_IDS[tg], N, K, x, nw, 1e-5, None)
    u, _ = _matvec_ref(wu, TYPE_IDS[tu], N, K, x, nw, 1e-5, None)
    ref = g / (1 + np.exp(-g)) * u
    assert np.abs(y - ref).max() / np.sqrt((ref ** 2).mean()) < 2e-3


def test_rmsnorm_vs_oracle(native_lib):
    from oracle.llama_ref import rmsnorm
    rng = np.random.default_rng(5)
    for K, rows in ((512, 1), (4096, 7), (2048, 0)):
        x = rng.standard_normal((max(rows, 1), K)).astype(np.float32) * 3
        w = rng.uniform(0.5, 1.5, size=K).astype(np.float32)
        y = np.zeros_like(x)
        lib_check(native_lib, native_lib.b2_op_rmsnorm(0, ptr(x), ptr(w), 1e-5, K, rows, ptr(y)))
        if rows:
            assert np.allclose(y, rmsnorm(x, w, 1e-5), rtol=1e-5, atol=1e-6)


def _attn_ref(q, k, v, n_head, n_kv, hd):
    T = k.shape[0]
    G = n_head // n_kv
    qf = q.reshape(n_kv, G, hd).astype(np.float64)
    kf = k.astype(np.float64)      # [T, n_kv, hd]
    vf = v.astype(np.float64)
    s = np.einsum("hgd,thd->hgt", qf, kf) / np.sqrt(hd)
    s = s - s.max(-1, keepdims=True)
    p = np.exp(s)
    p /= p.sum(-1, keepdims=True)
    return np.einsum("hgt,thd->hgd", p, vf).reshape(-1)


@pytest.mark.parametrize("n_head,n_kv,hd", [(8, 2, 64), (4, 2, 128), (32, 8, 128), (32, 4, 64), (4, 4, 128), (8, 1, 128)])
@pytest.mark.parametrize("T", [1, 15, 64, 65, 200, 1000])
def test_attention_decode_vs_oracle(native_lib, n_head, n_kv, hd, T):
    rng = np.random.default_rng(T * 7 + n_head)
    q = rng.standard_normal(n_head * hd).astype(np.float32)
    k = (rng.standard_normal((T, n_kv, hd)) * 0.7).astype(np.float16)
    v = rng.standard_normal((T, n_kv, hd)).astype(np.float16)
    ref = _attn_ref(q, k.astype(np.float32), v.astype(np.float32), n_head, n_kv, hd)
    for bt, n_splits in ((16, 1), (16, 5), (64, 18), (64, 8), (16, 4), (64, 2)):
        nblk = (T + bt - 1) // bt
        total = nblk + 3
        blocks = rng.permutation(total)[:nblk].astype(np.int32)      # scattered pages
        out = np.zeros(n_head * hd, np.float32)
        rc = native_lib.b2_op_attention_decode(0, ptr(q), ptr(k.view(np.uint16)), ptr(v.view(np.uint16)), T, n_head,
                                               n_kv, hd, bt, ptr(blocks), total, n_splits, ptr(out))
        lib_check(native_lib, rc)
        # float32 accumulation over f16 K/V: 1e-4 absolute on O(1) outputs
        assert np.abs(out - ref).max() < 2e-4, (n_head, n_kv, hd, T, bt, n_splits, np.abs(out - ref).max())


@pytest.mark.parametrize("n_head,n_kv,hd", [(8, 2, 64), (4, 2, 128), (32, 8, 128), (32, 4, 64), (4, 4, 128), (8, 1, 128)])
@pytest.mark.parametrize("T", [1, 15, 64, 65, 200, 1000, 3000])
def test_attention_decode_tensor_core_vs_oracle(native_lib, n_head, n_kv, hd, T):
    """attention_mma.cu (negative n_splits selects it): q and the probabilities are rounded to f16 for the MMA, so the
    reference uses the f16-rounded q and the bar is 2e-3 absolute on O(1) outputs (f16 P: 2^-11 relative per term)."""
    rng = np.random.default_rng(T * 11 + n_head)
    q = rng.standard_normal(n_head * hd).astype(np.float32)
    k = (rng.standard_normal((T, n_kv, hd)) * 0.7).astype(np.float16)
    v = rng.standard_normal((T, n_kv, hd)).astype(np.float16)
    ref = _attn_ref(q.astype(np.float16).astype(np.float32), k.astype(np.float32), v.astype(np.float32), n_head, n_kv, hd)
    for bt, n_splits in ((16, 1), (16, 5), (64, 32), (64, 8), (32, 3)):
        nblk = (T + bt - 1) // bt
        total = nblk + 3
        blocks = rng.permutation(total)[:nblk].astype(np.int32)
        out = np.zeros(n_head * hd, np.float32)
        rc = native_lib.b2_op_attention_decode(0, ptr(q), ptr(k.view(np.uint16)), ptr(v.view(np.uint16)), T, n_head,
                                               n_kv, hd, bt, ptr(blocks), total, -n_splits, ptr(out))
        lib_check(native_lib, rc)
        assert np.abs(out - ref).max() < 2e-3, (n_head, n_kv, hd, T, bt, n_splits, np.abs(out - ref).max())


def test_attention_rejects_bad_config(native_lib):
    q = np.zeros(96 * 2, np.float32)
    k = np.zeros((4, 2, 96), np.uint16)
    blocks = np.zeros(1, np.int32)
    out = np.zeros(96 * 2, np.float32)
    assert native_lib.b2_op_attention_decode(0, ptr(q), ptr(k), ptr(k), 4, 2, 2, 96, 16, ptr(blocks), 1, 1, ptr(out)) != 0


def _sample(native_lib, logits, hist, u, **kw):
    from llama_cpp._cabi import B2Sampling
    sp = B2Sampling(kw.get("temperature", 0.8), kw.get("top_p", 0.95), kw.get("min_p", 0.05),
                    kw.get("repeat_penalty", 1.1), kw.get("frequency_penalty", 0.0), kw.get("presence_penalty", 0.0),
                    kw.get("top_k", 40), kw.get("penalty_last_n", 64))
    h = np.asarray(hist, dtype=np.int32)
    tok = ctypes.c_int32(-1)
    rc = native_lib.b2_op_sample(0, ptr(logits), len(logits), ctypes.byref(sp), ptr(h) if len(h) else None, len(h),
                                 ctypes.c_double(u), ctypes.byref(tok))
    lib_check(native_lib, rc)
    return tok.value


@pytest.mark.parametrize("V", [1024, 32000, 128256])
def test_sampler_vs_oracle(native_lib, V):
    from oracle import sampler_ref as sr
    rng = np.random.default_rng(V)
    n_checked = 0
    for trial in range(12):
        logits = (rng.standard_normal(V) * rng.choice([0.5, 2.0, 6.0])).astype(np.float32)
        hist = rng.integers(0, V, size=rng.integers(0, 65)).tolist()
        if hist:
            hist[-1] = hist[0]          # a repeated token
        for kw in (dict(temperature=0.0, repeat_penalty=1.0),
                   dict(temperature=0.0),
                   dict(temperature=1.2, top_p=0.9, frequency_penalty=0.7, presence_penalty=0.8),   # api.py:59-62
                   dict(temperature=0.7, top_k=5, top_p=1.0, min_p=0.0),
                   dict(temperature=1.0, top_k=1000, top_p=0.5, min_p=0.2)):
            u = float(rng.uniform())
            ref_kw = dict(temperature=0.8, top_k=40, top_p=0.95, min_p=0.05, repeat_penalty=1.1,
                          frequency_penalty=0.0, presence_penalty=0.0, penalty_last_n=64)
            ref_kw.update(kw)
            want = sr.sample(logits, hist, u, **ref_kw)
            got = _sample(native_lib, logits, hist, u, **kw)
            if got != want and ref_kw["temperature"] > 0:
                # a draw that lands within float rounding of a cumulative-probability edge may
                # legitimately pick the neighbour: accept only if u is within 1e-6 of that edge
                l = sr.apply_penalties(logits, hist, 64, ref_kw["repeat_penalty"], ref_kw["frequency_penalty"],
                                       ref_kw["presence_penalty"])
                ids, p = sr.candidates_after_filters(l, ref_kw["top_k"], ref_kw["top_p"], ref_kw["min_p"],
                                                     ref_kw["temperature"])
                cp = np.cumsum(p.astype(np.float64) / p.astype(np.float64).sum())
                assert got in ids.tolist() and np.abs(cp - u).min() < 1e-6, (V, trial, kw, got, want)
            else:
                assert got == want, (V, trial, kw, got, want)
            n_checked += 1
    assert n_checked == 60


def test_sampler_greedy_ties_take_lowest_id(native_lib):
    logits = np.full(5000, -1.0, np.float32)
    logits[[4321, 77, 900]] = 3.5
    assert _sample(native_lib, logits, [], 0.5, temperature=0.0, repeat_penalty=1.0) == 77


def test_sampler_topk_ties_at_threshold(native_lib):
    # 100 equal logits compete for the last top-k places: the lowest ids must be kept
    from oracle import sampler_ref as sr
    logits = np.full(4096, -5.0, np.float32)
    logits[10:110] = 1.0
    logits[3000] = 2.0
    for u in (0.01, 0.5, 0.99):
        want = sr.sample(logits, [], u, temperature=1.0, top_k=8, top_p=1.0, min_p=0.0, repeat_penalty=1.0)
        got = _sample(native_lib, logits, [], u, temperature=1.0, top_k=8, top_p=1.0, min_p=0.0, repeat_penalty=1.0)
        assert got == want and (got == 3000 or 10 <= got < 17)


def test_sampler_rejects_unsupported_top_k(native_lib):
    logits = np.zeros(100, np.float32)
    from llama_cpp._cabi import B2Sampling
    sp = B2Sampling(1.0, 1.0, 0.0, 1.0, 0.0, 0.0, 5000, 64)
    tok = ctypes.c_int32()
    assert native_lib.b2_op_sample(0, ptr(logits), 100, ctypes.byref(sp), None, 0, ctypes.c_double(0.5), ctypes.byref(tok)) != 0


# ---------------------------------------------------------------------------------------------------
# tensor-core GEMM (prefill / batched decode): y[t][n] = sum_k f16(x[t][k]) * W[n][k]
def _gemm_case(native_lib, tname, K, N, T, accumulate=False, n_splits=0):
    from llama_cpp.synth import random_blocks
    from oracle import ggml_blocks as gb
    t = TYPE_IDS[tname]
    be = gb.BLOCK[t][0]
    rng = np.random.default_rng(K * 7 + N * 3 + T)
    w = random_blocks(tname, N * K // be, 0.02, seed=K * 13 + N + T).reshape(-1)
    x = (rng.standard_normal((T, K)) * rng.uniform(0.2, 2.0, size=K)).astype(np.float32)
    y_in = rng.standard_normal((T, N)).astype(np.float32) if accumulate else None
    y = np.zeros((T, N), np.float32)
    rc = native_lib.b2_op_gemm(0, t, ptr(w), N, K, ptr(x), T, ptr(y_in) if accumulate else None, ptr(y), n_splits)
    lib_check(native_lib, rc)
    W = gb.dequantize(w, t).reshape(N, K).astype(np.float64)
    ref = x.astype(np.float16).astype(np.float64) @ W.T
    rms = np.sqrt((ref ** 2).mean())
    if accumulate:
        ref = ref + y_in
    return np.abs(y - ref).max() / rms


@pytest.mark.parametrize("tname", QTYPES)
@pytest.mark.parametrize("K,N,T", [(256, 128, 16), (512, 256, 1), (1024, 300, 37), (4096, 1024, 128),
                                   (2048, 128, 256), (2048, 384, 300), (14336, 256, 512)])
def test_gemm_vs_oracle(native_lib, tname, K, N, T):
    # weights are rounded to f16 once (2^-11 relative each), products accumulate in float32 on the tensor core
    err = _gemm_case(native_lib, tname, K, N, T)
    assert err < 2e-3, (tname, K, N, T, err)


def test_gemm_accumulates_into_residual(native_lib):
    assert _gemm_case(native_lib, "Q4_K", 1024, 256, 48, accumulate=True) < 2e-3


@pytest.mark.parametrize("n_splits", [1, 2, 3, 16])
def test_gemm_k_splits_add_up(native_lib, n_splits):
    # K ranges computed by separate CTAs (uneven when K/64 is not a multiple) and added in a fixed order
    assert _gemm_case(native_lib, "Q6_K", 1280, 200, 40, n_splits=n_splits) < 2e-3
    assert _gemm_case(native_lib, "Q8_0", 4096, 128, 320, n_splits=n_splits) < 2e-3


def test_gemm_rejects_bad_shapes(native_lib):
    w = np.zeros(144 * 128, np.uint8)
    x = np.zeros((4, 256), np.float32)
    y = np.zeros((4, 128), np.float32)
    assert native_lib.b2_op_gemm(0, TYPE_IDS["Q4_K"], ptr(w), 128, 128, ptr(x), 4, None, ptr(y), 0) != 0     # K % 256
    assert native_lib.b2_op_gemm(0, TYPE_IDS["Q4_K"], ptr(w), 128, 256, ptr(x), 513, None, ptr(y), 0) != 0   # too many tokens
    assert native_lib.b2_op_gemm(0, TYPE_IDS["F32"], ptr(w), 128, 256, ptr(x), 4, None, ptr(y), 0) != 0      # not a block format
