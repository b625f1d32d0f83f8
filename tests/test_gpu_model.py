"""End-to-end parity of the native engine against the CPU oracle on seeded synthetic models
(GPU only). Bars from BASELINE.json: per-token logits max-abs <= 1e-2; greedy tokens identical
(checked where the oracle's top-2 margin exceeds 10x the logit tolerance)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LOGIT_TOL = 1e-2
MODELS = [("tiny", "Q4_K_M", 1), ("tiny128", "Q4_0", 2), ("tiny", "Q8_0", 3), ("tiny128", "Q5_K", 4), ("tiny", "Q6_K", 5),
          # block formats the GEMM path dequantises element by element
          ("tiny", "Q4_1", 6), ("tiny128", "Q5_0", 7), ("tiny", "Q5_1", 8), ("tiny128", "Q2_K", 9), ("tiny", "Q3_K", 10),
          ("tiny", "Q2_K_mix", 11)]          # one launch mixes formats of both kernel instantiations
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "tiny_model.npz"))
TOKENS = [int(t) for t in GOLD["tokens"]]


def _model(tmp_models, arch, ftype, seed):
    from llama_cpp import synth
    path = os.path.join(tmp_models, f"{arch}_{ftype}_{seed}.gguf")
    if not os.path.exists(path):
        synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=0.1)
    return path


def _engine(path, flags=0, max_seqs=2, n_ctx=256, bt=16):
    from llama_cpp.engine import Engine
    from llama_cpp.gguf_file import GGUFFile
    return Engine(GGUFFile(path), n_ctx=n_ctx, max_seqs=max_seqs, kv_block_tokens=bt, flags=flags)


def _greedy():
    from llama_cpp.engine import SamplingSettings
    return SamplingSettings(temperature=0.0, repeat_penalty=1.0)


@pytest.mark.parametrize("arch,ftype,seed", MODELS)
def test_logits_prefill_and_teacher_forced_decode(native_lib, tmp_models, arch, ftype, seed):
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, arch, ftype, seed)
    ref = RefModel(path).forward(TOKENS)                       # [T, V] float32 oracle
    eng = _engine(path)
    eng.ensure_pages(0, len(TOKENS))
    eng.seq_begin(0, 0, [], _greedy())
    eng.prefill(0, TOKENS[:16])
    worst = np.abs(eng.logits(0) - ref[15]).max()
    for i in range(16, len(TOKENS)):
        eng.set_token(0, TOKENS[i])
        eng.forward_pending(0)
        worst = max(worst, np.abs(eng.logits(0) - ref[i]).max())
    assert worst <= LOGIT_TOL, (arch, ftype, worst)
    key = f"{arch}_{ftype}_{seed}_logits"
    if key in GOLD:   # independent implementation (transformers, no f16 KV rounding): 2x the bar
        assert np.abs(eng.logits(0) - GOLD[key][-1]).max() <= 2 * LOGIT_TOL
    eng.close()


@pytest.mark.parametrize("arch,ftype,seed", MODELS[:3])
def test_greedy_sequence_matches_oracle(native_lib, tmp_models, arch, ftype, seed):
    """128 greedy tokens. The oracle runs teacher-forced over the ENGINE's tokens, so after a legitimate near-tie flip
    the comparison re-synchronises and keeps checking the remaining steps instead of stopping at the first difference:
    every step whose oracle top-2 margin is >= 10x the logit tolerance must carry the oracle's arg-max."""
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, arch, ftype, seed)
    n_new = 128
    prompt = TOKENS[:12]
    eng = _engine(path)
    eng.ensure_pages(0, len(prompt) + n_new + 1)
    eng.seq_begin(0, 0, [], _greedy())
    eng.prefill(0, prompt)
    got = [eng.sample(0)]
    for _ in range(n_new - 1):
        got.append(eng.decode([0])[0])
    eng.close()
    want = RefModel(path).forward(prompt + got[:-1])[len(prompt) - 1:]     # logits behind each generated token
    top2 = np.partition(want, -2, axis=1)[:, -2:]
    margin = top2[:, 1] - top2[:, 0]
    flips = [i for i in range(n_new) if int(want[i].argmax()) != got[i]]
    for i in flips:
        assert margin[i] < 10 * LOGIT_TOL, f"step {i}: engine {got[i]} != oracle {int(want[i].argmax())} with margin {margin[i]}"
    checked = int((margin >= 10 * LOGIT_TOL).sum())
    assert checked >= n_new // 2, f"only {checked} of {n_new} steps have a clear margin: pick another seed"


def test_graph_pdl_and_plain_launches_agree_bitwise(native_lib, tmp_models):
    from llama_cpp import _cabi
    path = _model(tmp_models, "tiny", "Q4_K_M", 1)
    outs = []
    for flags in (0, _cabi.B2_FLAG_NO_PDL, _cabi.B2_FLAG_NO_GRAPH | _cabi.B2_FLAG_NO_PDL, _cabi.B2_FLAG_NO_GRAPH):
        eng = _engine(path, flags=flags)
        eng.ensure_pages(0, 64)
        eng.seq_begin(0, 0, [], _greedy())
        eng.prefill(0, TOKENS[:10])
        toks = [eng.sample(0)] + [eng.decode([0])[0] for _ in range(20)]
        outs.append((toks, eng.logits(0).copy(), eng.mode()))
        eng.close()
    for toks, lg, mode in outs[1:]:
        assert toks == outs[0][0], mode
        assert np.array_equal(lg, outs[0][1]), mode


def test_two_slots_do_not_interfere(native_lib, tmp_models):
    path = _model(tmp_models, "tiny", "Q4_K_M", 1)
    eng = _engine(path, max_seqs=3)
    pa, pb = TOKENS[:9], TOKENS[5:20]

    def solo(slot, prompt, n):
        eng.ensure_pages(slot, len(prompt) + n + 1)
        eng.seq_begin(slot, 0, [], _greedy())
        eng.prefill(slot, prompt)
        out = [eng.sample(slot)] + [eng.decode([slot])[0] for _ in range(n - 1)]
        eng.release(slot)
        return out

    a_alone, b_alone = solo(0, pa, 16), solo(0, pb, 16)
    for s, pr in ((1, pa), (2, pb)):
        eng.ensure_pages(s, len(pr) + 17)
        eng.seq_begin(s, 0, [], _greedy())
        eng.prefill(s, pr)
    a, b = [eng.sample(1)], [eng.sample(2)]
    for _ in range(15):
        ta, tb = eng.decode([1, 2])
        a.append(ta)
        b.append(tb)
    assert a == a_alone and b == b_alone
    eng.close()


def test_sampled_decode_matches_oracle_sampler(native_lib, tmp_models):
    """temperature 1.2 / top_p 0.9 / penalties as in api.py:59-62, same uniforms on both sides."""
    from llama_cpp.engine import SamplingSettings
    from oracle import sampler_ref as sr
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, "tiny", "Q4_K_M", 1)
    prompt = TOKENS[:10]
    st = SamplingSettings(temperature=1.2, top_p=0.9, frequency_penalty=0.7, presence_penalty=0.8)
    u = sr.MT19937Uniforms(1234).take(64)
    eng = _engine(path)
    eng.ensure_pages(0, 80)
    eng.seq_begin(0, 0, [], st, uniforms=u)
    eng.prefill(0, prompt)
    got = [eng.sample(0)] + [eng.decode([0])[0] for _ in range(23)]
    # oracle, teacher-forced on the engine's tokens so one near-tie cannot cascade
    ref = RefModel(path)
    lg = ref.forward(prompt, last_only=True)[0]
    hist = list(prompt)
    agree = 0
    for i, t in enumerate(got):
        want = sr.sample(lg, hist, float(u[i]), temperature=1.2, top_k=40, top_p=0.9, min_p=0.05, repeat_penalty=1.1,
                         frequency_penalty=0.7, presence_penalty=0.8)
        agree += int(want == t)
        hist.append(t)
        lg = ref.forward([t], last_only=True)[0]
    assert agree >= len(got) - 1, (agree, len(got))     # at most one boundary-rounding disagreement
    eng.close()


def test_engine_rejects_bad_input(native_lib, tmp_models):
    from llama_cpp._cabi import B2Error
    path = _model(tmp_models, "tiny", "Q4_K_M", 1)
    eng = _engine(path, n_ctx=32)
    eng.ensure_pages(0, 32)
    eng.seq_begin(0, 0, [], _greedy())
    with pytest.raises(B2Error):
        eng.prefill(0, list(range(40)))          # longer than n_ctx
    with pytest.raises(B2Error):
        eng.prefill(0, [10 ** 6])                # token id out of range
    with pytest.raises(B2Error):
        eng.prefill(5, [1])                      # slot out of range
    eng.close()


@pytest.mark.parametrize("arch,ftype,seed", MODELS)
def test_tensor_core_prefill_matches_oracle(native_lib, tmp_models, arch, ftype, seed):
    """Prompt chunks of >= 8 tokens go through the tcgen05 GEMM path (f16 activations, f32 accumulate): logits
    of the last prompt token against the oracle, for one whole-prompt call, for a prompt split into two calls
    (the second chunk attends to the first through the paged cache) and for decode continuing after it."""
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, arch, ftype, seed)
    ref = RefModel(path).forward(TOKENS)
    n = len(TOKENS)
    for split in (None, 11):
        eng = _engine(path)
        eng.ensure_pages(0, n)
        eng.seq_begin(0, 0, [], _greedy())
        k = n - 4
        if split is None:
            eng.prefill(0, TOKENS[:k])
        else:
            eng.prefill(0, TOKENS[:split])
            assert np.abs(eng.logits(0) - ref[split - 1]).max() <= LOGIT_TOL
            eng.prefill(0, TOKENS[split:k])
        worst = np.abs(eng.logits(0) - ref[k - 1]).max()
        for i in range(k, n):                     # teacher-forced decode on top of the GEMM-written KV cache
            eng.set_token(0, TOKENS[i])
            eng.forward_pending(0)
            worst = max(worst, np.abs(eng.logits(0) - ref[i]).max())
        assert worst <= LOGIT_TOL, (arch, ftype, split, worst)
        eng.close()


def test_tensor_core_prefill_agrees_with_token_by_token(native_lib, tmp_models, monkeypatch):
    """Same prompt through the GEMM path and through the mat-vec path (B2_PREFILL_MIN above the prompt
    length): same greedy continuation, logits within the tolerance; a 700-token prompt spans two 512-token chunks."""
    path = _model(tmp_models, "tiny128", "Q4_K_M", 9)
    rng = np.random.default_rng(5)
    prompt = [int(t) for t in rng.integers(3, 700, size=700)]
    outs = []
    for by_token in (False, True):
        if by_token:
            monkeypatch.setenv("B2_PREFILL_MIN", "100000")
        else:
            monkeypatch.delenv("B2_PREFILL_MIN", raising=False)
        eng = _engine(path, n_ctx=1024, bt=64)
        eng.ensure_pages(0, 800)
        eng.seq_begin(0, 0, [], _greedy())
        eng.prefill(0, prompt)
        lg = eng.logits(0).copy()
        toks = [eng.sample(0)] + [eng.decode([0])[0] for _ in range(8)]
        outs.append((toks, lg))
        eng.close()
    assert np.abs(outs[0][1] - outs[1][1]).max() <= LOGIT_TOL
    assert outs[0][0] == outs[1][0]


@pytest.mark.parametrize("attn_splits", [0, 1])      # 1: one split per sequence (what 32-sequence steps use)
@pytest.mark.parametrize("arch,ftype,seed", [("tiny", "Q4_K_M", 1), ("tiny128", "Q8_0", 6)])
def test_batched_decode_matches_oracle_and_serial_path(native_lib, tmp_models, arch, ftype, seed, attn_splits, monkeypatch):
    """>= 3 sequences per b2_decode call step together through the tensor-core GEMM (weights read once per
    step), each at its own position with its own pages and sampler state. Greedy continuations must equal the
    one-sequence-at-a-time path wherever the oracle's top-2 margin is clear, logits within the tolerance."""
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, arch, ftype, seed)
    ref = RefModel(path)
    rng = np.random.default_rng(11)
    prompts = [[int(t) for t in rng.integers(3, 700, size=n)] for n in (5, 9, 17, 30, 12, 3)]
    n_new = 10
    runs = {}
    for mode in ("serial", "batched"):
        monkeypatch.setenv("B2_BATCH_MIN", "100" if mode == "serial" else "3")
        if attn_splits:
            monkeypatch.setenv("B2_BATCH_ATTN_SPLITS", str(attn_splits))
        eng = _engine(path, max_seqs=8, n_ctx=128)
        for s, pr in enumerate(prompts):
            eng.ensure_pages(s, len(pr) + n_new + 2)
            eng.seq_begin(s, 0, [], _greedy())
            eng.prefill(s, pr)
        seqs = [[eng.sample(s)] for s in range(len(prompts))]
        for _ in range(n_new - 1):
            for s, t in enumerate(eng.decode(list(range(len(prompts))))):
                seqs[s].append(t)
        runs[mode] = (seqs, [eng.logits(s).copy() for s in range(len(prompts))])
        eng.close()
    for s, pr in enumerate(prompts):
        serial, batched = runs["serial"][0][s], runs["batched"][0][s]
        # oracle logits along the serial continuation: tokens must agree wherever the margin is clear
        full = pr + serial
        ref.reset()                      # the oracle appends to its own KV cache: one sequence per reset
        lg = ref.forward(full)
        for i in range(n_new):
            row = lg[len(pr) - 1 + i]
            top2 = np.sort(row)[-2:]
            if batched[:i] != serial[:i]:
                break
            if top2[1] - top2[0] >= 10 * LOGIT_TOL:
                assert batched[i] == serial[i] == int(np.argmax(row)), (s, i)
        if batched == serial:
            assert np.abs(runs["batched"][1][s] - runs["serial"][1][s]).max() <= LOGIT_TOL
            assert np.abs(runs["batched"][1][s] - lg[len(full) - 2]).max() <= LOGIT_TOL


def test_batched_decode_rejects_duplicate_slots(native_lib, tmp_models):
    from llama_cpp._cabi import B2Error
    path = _model(tmp_models, "tiny", "Q4_K_M", 1)
    eng = _engine(path, max_seqs=4)
    for s in range(2):
        eng.ensure_pages(s, 16)
        eng.seq_begin(s, 0, [], _greedy())
        eng.prefill(s, TOKENS[:4])
        eng.sample(s)
    with pytest.raises(B2Error):
        eng.decode([0, 1, 0])
    eng.close()


def test_long_context_uses_more_attention_splits_and_matches_oracle(native_lib, tmp_models):
    """n_ctx > 2048 switches flash-decode to 16 splits merged through global scratch, and a 2600-token prompt spans
    six 512-token GEMM chunks whose causal attention reads the earlier chunks from the paged cache."""
    from oracle.llama_ref import RefModel
    path = _model(tmp_models, "tiny128", "Q4_K_M", 9)
    rng = np.random.default_rng(21)
    prompt = [int(t) for t in rng.integers(3, 700, size=2600)]
    ref = RefModel(path)
    lg = ref.forward(prompt, last_only=True)[-1]
    eng = _engine(path, n_ctx=5000, bt=64, max_seqs=1)
    assert eng.mode()["attn_splits"] == 16
    eng.ensure_pages(0, 2700)
    eng.seq_begin(0, 0, [], _greedy())
    eng.prefill(0, prompt)
    assert np.abs(eng.logits(0) - lg).max() <= LOGIT_TOL
    tok = eng.sample(0)
    for _ in range(3):                       # decode steps attend to all 2600+ cached positions
        nxt = ref.forward([tok], last_only=True)[-1]
        t2 = eng.decode([0])[0]
        assert np.abs(eng.logits(0) - nxt).max() <= LOGIT_TOL
        top2 = np.sort(nxt)[-2:]
        if top2[1] - top2[0] >= 10 * LOGIT_TOL:
            assert t2 == int(np.argmax(nxt))
        tok = t2
    eng.close()
