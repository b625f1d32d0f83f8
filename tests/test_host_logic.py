"""Host-side logic on the CPU with a fake engine: tokenizers, chat templates, the drop-in Llama
surface, the continuous-batching scheduler (prefix reuse, run-ahead + rewind, stop handling)."""
import os
import threading

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from fake_engine import FakeEngine


@pytest.fixture(scope="module")
def tiny_path(tmp_path_factory):
    from llama_cpp import synth
    p = str(tmp_path_factory.mktemp("host") / "tiny.gguf")
    synth.write_synthetic_model(p, "tiny", "Q4_K_M", seed=1)
    return p


@pytest.fixture()
def llm(tiny_path, monkeypatch):
    import llama_cpp.llama as L
    monkeypatch.setattr(L, "Engine", FakeEngine)
    FakeEngine.delay, FakeEngine.eos_after, FakeEngine.fail_with = 0.0, None, None
    m = L.Llama(model_path=tiny_path, n_gpu_layers=-1, n_ctx=128, n_seq_max=4, seed=3, verbose=False)
    FakeEngine.eos_token = m._vocab.eot_id
    yield m
    m.close()


# ---------------------------------------------------------------- tokenizers
def test_spm_tokenizer_matches_sentencepiece():
    spm = pytest.importorskip("sentencepiece")
    mc = pytest.importorskip("mistral_common")
    from llama_cpp import synth
    from llama_cpp.tokenizer import Vocab
    v = Vocab(synth.synth_spm_vocab(32000))
    sp = spm.SentencePieceProcessor(model_file=os.path.join(os.path.dirname(mc.__file__), "data", "tokenizer.model.v1"))
    for text in ["Hello world, this is a test.", "  two  spaces", "naïve café €5 — 日本語", "a\nb\n\nc", "", "x" * 50]:
        mine = v.tokenize(text.encode(), add_bos=False)
        assert mine == sp.encode(text), text
        if text:
            assert v.detokenize([v.bos_id] + mine).decode() == text
    assert v.tokenize(b"hi", add_bos=True)[0] == v.bos_id == 1


def test_bpe_tokenizer_round_trip_and_specials(tiny_path):
    from llama_cpp.gguf_file import GGUFFile
    from llama_cpp.tokenizer import Vocab
    v = Vocab(GGUFFile(tiny_path).metadata)
    for text in ["hello world", "the time is now! 123", "tabs\tand\nnewlines", "unicodé ✓"]:
        ids = v.tokenize(text.encode(), add_bos=False)
        assert v.detokenize(ids).decode() == text
    ids = v.tokenize(b"<|start_header_id|>user<|end_header_id|>\n\nhi<|eot_id|>", add_bos=False, special=True)
    assert ids[0] == v.special[b"<|start_header_id|>"] and ids[-1] == v.eot_id
    assert v.detokenize(ids) == b"user\n\nhi"                       # control tokens render as nothing
    plain = v.tokenize(b"<|eot_id|>", add_bos=False, special=False)   # not parsed: plain text
    assert v.eot_id not in plain and v.detokenize(plain) == b"<|eot_id|>"
    assert v.is_eog(v.eot_id) and v.is_eog(v.eos_id) and not v.is_eog(5)
    assert v.tokenize(b"", add_bos=True) == [v.bos_id]               # like upstream: BOS unless the file says add_bos_token = false
    # ignore_merges (llama-bpe): a pre-token that is itself in the vocabulary is emitted whole
    whole = next(t for t in v.tokens[300:500] if t.startswith(b"\xc4\xa0") and len(t) > 4)   # "Ġword" entries of the synthetic vocabulary
    word = whole.decode().replace("\u0120", " ")
    assert v.tokenize(word.encode(), add_bos=False) == [v.token_to_id[whole]]


@settings(max_examples=60, deadline=None)
@given(st.text(alphabet=st.characters(blacklist_categories=("Cs",)), max_size=40))
def test_bpe_round_trip_property(tiny_path_holder, text):
    v = tiny_path_holder
    assert v.detokenize(v.tokenize(text.encode("utf-8"), add_bos=False)).decode("utf-8") == text


@pytest.fixture(scope="module")
def tiny_path_holder(tiny_path):
    from llama_cpp.gguf_file import GGUFFile
    from llama_cpp.tokenizer import Vocab
    return Vocab(GGUFFile(tiny_path).metadata)


# ---------------------------------------------------------------- chat templates
def test_llama3_format_drops_unknown_roles():
    from llama_cpp import chat_format as cf
    r = cf.format_llama3([{"role": "user", "content": "hi"}, {"role": "bot", "content": "dropped"},
                          {"role": "system", "content": "sys"}])
    assert r.prompt == ("<|begin_of_text|><|start_header_id|>user<|end_header_id|>\n\nhi<|eot_id|>"
                        "<|start_header_id|>system<|end_header_id|>\n\nsys<|eot_id|>"
                        "<|start_header_id|>assistant<|end_header_id|>\n\n")
    assert r.stop == "<|eot_id|>" and not r.added_special


def test_other_formats_and_guessing():
    from llama_cpp import chat_format as cf
    msgs = [{"role": "system", "content": "S"}, {"role": "user", "content": "U"}, {"role": "assistant", "content": "A"},
            {"role": "user", "content": "U2"}]
    assert cf.format_chatml(msgs).prompt == ("<|im_start|>system\nS<|im_end|>\n<|im_start|>user\nU<|im_end|>\n"
                                             "<|im_start|>assistant\nA<|im_end|>\n<|im_start|>user\nU2<|im_end|>\n<|im_start|>assistant\n")
    assert cf.format_mistral_instruct(msgs).prompt == "<s>[INST] U [/INST]A</s>[INST] U2 [/INST]"
    assert cf.format_llama2(msgs).prompt.startswith("<s>[INST] <<SYS>>\nS\n<</SYS>>")
    assert cf.guess_chat_format_from_gguf_metadata({"tokenizer.chat_template": cf.LLAMA3_INSTRUCT_CHAT_TEMPLATE}) == "llama-3"
    assert cf.guess_chat_format_from_gguf_metadata({"tokenizer.chat_template": cf.CHATML_CHAT_TEMPLATE}) == "chatml"
    assert cf.guess_chat_format_from_gguf_metadata({"tokenizer.chat_template": "{{ custom }}"}) is None
    assert cf.guess_chat_format_from_gguf_metadata({}) is None
    j = cf.Jinja2ChatFormatter(cf.CHATML_CHAT_TEMPLATE, "</s>", "<s>")([{"role": "user", "content": "x"}])
    assert j.prompt == "<|im_start|>user\nx<|im_end|>\n<|im_start|>assistant\n" and j.added_special and j.stop == ["</s>"]


# ---------------------------------------------------------------- Llama surface on the fake engine
def test_chat_completion_shape_and_usage(llm):
    msgs = [{"role": "user", "content": "hello world"}]
    out = llm.create_chat_completion(messages=msgs, stream=False, temperature=1.2, top_p=0.9,
                                     frequency_penalty=0.7, presence_penalty=0.8, max_tokens=10)
    assert out["object"] == "chat.completion" and out["choices"][0]["finish_reason"] == "length"
    assert out["usage"]["completion_tokens"] == 10 and out["model"] == llm.model_path
    assert isinstance(out["choices"][0]["message"]["content"], str) and msgs == [{"role": "user", "content": "hello world"}]
    assert llm.chat_format == "llama-3"


def test_context_window_error_and_max_tokens_default(llm):
    with pytest.raises(ValueError, match=r"Requested tokens \(200\) exceed context window of 128"):
        llm.create_completion(list(range(1, 201)))
    out = llm.create_completion([5, 6, 7], max_tokens=None)          # None -> fill the context
    assert out["usage"]["total_tokens"] == 128 and out["choices"][0]["finish_reason"] == "length"


def test_eog_token_stops_generation(llm):
    FakeEngine.eos_after = 5
    out = llm.create_completion([9, 10, 11], max_tokens=40, temperature=0.0)
    assert out["choices"][0]["finish_reason"] == "stop" and out["usage"]["completion_tokens"] == 5
    assert llm._engine.calls["rewind"] >= 0


def test_stop_string_cuts_text_and_rewinds(llm):
    full = llm.create_completion([9, 10, 11], max_tokens=30, temperature=0.0)["choices"][0]["text"]
    assert len(full) > 8
    stop = full[5:8]
    llm.reset()
    out = llm.create_completion([9, 10, 11], max_tokens=30, temperature=0.0, stop=[stop])
    assert out["choices"][0]["finish_reason"] == "stop"
    assert out["choices"][0]["text"] == full[:full.index(stop)]
    assert llm._engine.calls["rewind"] >= 1                          # run-ahead tokens past the stop were discarded
    # after the rewind the cached prefix is consistent: the same request again reuses it fully
    out2 = llm.create_completion([9, 10, 11], max_tokens=30, temperature=0.0, stop=[stop])
    assert out2["choices"][0]["text"] == out["choices"][0]["text"] and llm.last_timings["reused_tokens"] == 2


def test_prefix_reuse_skips_prefill(llm):
    prompt = list(range(20, 60))
    llm.create_completion(prompt, max_tokens=4, temperature=0.0)
    n0 = llm._engine.calls["prefill_tokens"]
    llm.create_completion(prompt + [7, 8], max_tokens=4, temperature=0.0)
    assert llm._engine.calls["prefill_tokens"] - n0 == 2 and llm.last_timings["reused_tokens"] == 40


def test_unsupported_options_raise(llm):
    for kw in (dict(top_k=0, temperature=1.0), dict(mirostat_mode=1), dict(logit_bias={"1": 2.0}), dict(typical_p=0.5),
               dict(logprobs=3)):
        with pytest.raises(NotImplementedError):
            llm.create_completion([1, 2], max_tokens=2, **kw)


def test_engine_error_reaches_caller(llm):
    FakeEngine.fail_with = RuntimeError("b200 engine call failed (code -2): boom")
    with pytest.raises(RuntimeError, match="boom"):
        llm.create_completion([1, 2, 3], max_tokens=2)
    FakeEngine.fail_with = None
    assert llm.create_completion([1, 2, 3], max_tokens=2)["usage"]["completion_tokens"] == 2   # scheduler survived


def test_streaming_chunks(llm):
    chunks = list(llm.create_chat_completion(messages=[{"role": "user", "content": "hi"}], stream=True, max_tokens=6))
    assert chunks[0]["choices"][0]["delta"] == {"role": "assistant"}
    assert chunks[-1]["choices"][0]["finish_reason"] in ("stop", "length")
    assert all(c["object"] == "chat.completion.chunk" for c in chunks)


# ---------------------------------------------------------------- scheduler: batching invariants
def test_concurrent_callers_are_batched_and_independent(llm):
    rng = np.random.default_rng(0)
    prompts = [[int(t) for t in rng.integers(1, 250, size=n)] for n in (5, 17, 9, 30, 12, 3, 21, 8)]
    serial = []
    for p in prompts:
        llm.reset()
        serial.append(llm.create_completion(p, max_tokens=12, temperature=0.0)["choices"][0]["text"])
    llm.reset()
    FakeEngine.delay = 0.002
    out = [None] * len(prompts)

    def work(i):
        out[i] = llm.create_completion(prompts[i], max_tokens=12, temperature=0.0)["choices"][0]["text"]

    th = [threading.Thread(target=work, args=(i,)) for i in range(len(prompts))]
    [t.start() for t in th]
    [t.join(60) for t in th]
    assert out == serial                                   # output independent of batch composition
    assert 2 <= llm._engine.calls["max_batch"] <= 4        # n_seq_max = 4 slots
    owned = [p for s in range(4) for p in llm._engine.pages.table(s)]
    assert len(owned) == len(set(owned))                   # no KV page owned twice


@settings(max_examples=25, deadline=None)
@given(st.lists(st.tuples(st.integers(1, 40), st.integers(1, 20)), min_size=1, max_size=6))
def test_scheduler_property_serial_equals_batched(tiny_path, reqs):
    import llama_cpp.llama as L
    orig = L.Engine
    L.Engine = FakeEngine
    FakeEngine.delay, FakeEngine.eos_after, FakeEngine.fail_with = 0.0, None, None
    try:
        m = L.Llama(model_path=tiny_path, n_gpu_layers=-1, n_ctx=96, n_seq_max=3, verbose=False)
        prompts = [[(7 * i + 3 * j) % 250 + 1 for j in range(n)] for i, (n, _) in enumerate(reqs)]
        want = []
        for p, (_, mt) in zip(prompts, reqs):
            m.reset()
            want.append(m.create_completion(p, max_tokens=mt, temperature=0.0)["usage"])
        m.reset()
        got = [None] * len(reqs)

        def work(i):
            got[i] = m.create_completion(prompts[i], max_tokens=reqs[i][1], temperature=0.0)["usage"]

        th = [threading.Thread(target=work, args=(i,)) for i in range(len(reqs))]
        [t.start() for t in th]
        [t.join(60) for t in th]
        assert got == want
        m.close()
    finally:
        L.Engine = orig
