"""The byte-level BPE path (llama-bpe pre-tokeniser, merges, ignore_merges) pinned against an independent implementation:
Hugging Face `tokenizers` configured the way Llama-3's tokenizer.json is (Split on the Llama-3 regex + ByteLevel, BPE with
ignore_merges). A small vocabulary is trained in the test itself (deterministic), exported the way a GGUF file carries it
(tokens, merges, token types) and both sides encode the same texts. Reference call sites: /root/reference/api.py:55-63
(create_chat_completion tokenises the rendered chat template)."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")]

tokenizers = pytest.importorskip("tokenizers")

CORPUS = [
    "The quick brown fox jumps over the lazy dog. The dog doesn't care; it's asleep!",
    "def tokenize(text: str) -> list[int]:\n    return [ord(c) for c in text]\n",
    "In 2024 the model processed 1234567 tokens at 548.6 tok/s -- 39% of the roofline.",
    "Grüße aus München! Привет, мир. こんにちは世界。 ¿Qué tal? naïve café",
    "user: hello there\nassistant: General Kenobi!   You're a bold one.\n\n\nuser: we'll see, I'd say",
    "    indented\tcode\twith\ttabs   and   runs   of   spaces    \r\n and CRLF",
] * 4

TEXTS = [
    "Hello world", " Hello  world ", "it's we'll I'd THEY'RE don't", "1234567890 12 345 6", "tok/s -- 39%!!! ???",
    "naïve café Привет こんにちは", "line one\nline two\n\n\nline five", "tabs\tand\r\nCRLF  ", "", " ", "a", "The The The dog dog",
    "def f(x):\n    return x**2  # comment", "emoji 🙂 and 👍🏽 mixed", "x" * 50, "word " * 20,
]


@pytest.fixture(scope="module")
def pair():
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers
    from llama_cpp.tokenizer import LLAMA3_PRE_REGEX, Vocab
    tok = Tokenizer(models.BPE(ignore_merges=True))
    tok.pre_tokenizer = pre_tokenizers.Sequence([
        pre_tokenizers.Split(Regex(LLAMA3_PRE_REGEX), behavior="isolated", invert=False),
        pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    specials = ["<|begin_of_text|>", "<|eot_id|>"]
    trainer = trainers.BpeTrainer(vocab_size=700, special_tokens=specials, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                  show_progress=False)
    tok.train_from_iterator(CORPUS, trainer)
    model = json.loads(tok.to_str())["model"]
    vocab = model["vocab"]
    tokens = [None] * len(vocab)
    for t, i in vocab.items():
        tokens[i] = t.encode("utf-8")
    merges = [m if isinstance(m, str) else " ".join(m) for m in model["merges"]]
    types = [3 if t.decode() in specials else 1 for t in tokens]        # CONTROL / NORMAL
    md = {"tokenizer.ggml.model": "gpt2", "tokenizer.ggml.pre": "llama-bpe", "tokenizer.ggml.tokens": tokens,
          "tokenizer.ggml.merges": merges, "tokenizer.ggml.token_type": types,
          "tokenizer.ggml.bos_token_id": vocab["<|begin_of_text|>"], "tokenizer.ggml.eos_token_id": vocab["<|eot_id|>"]}
    return tok, Vocab(md)


@pytest.mark.parametrize("text", TEXTS)
def test_bpe_ids_match_hf_tokenizers(pair, text):
    tok, vocab = pair
    want = tok.encode(text, add_special_tokens=False).ids
    got = vocab.tokenize(text.encode("utf-8"), add_bos=False, special=False)
    assert got == want, (text, got, want)
    if text:
        assert vocab.detokenize(got).decode("utf-8") == text


def test_special_tokens_split_and_bos(pair):
    tok, vocab = pair
    text = "<|begin_of_text|>hi there<|eot_id|>"
    want = tok.encode(text, add_special_tokens=False).ids
    assert vocab.tokenize(text.encode(), add_bos=False, special=True) == want
    assert vocab.tokenize(b"hi", add_bos=True, special=False)[0] == vocab.bos_id


def test_bpe_ids_match_hf_tokenizers_on_random_text(pair):
    hyp = pytest.importorskip("hypothesis")
    from hypothesis import given, settings, strategies as st
    tok, vocab = pair
    alphabet = st.sampled_from(list("abcdefghijklmnopqrstuvwxyzABCDEFG0123456789 \n\t\r'.,!?-_/()#*%é ñüßПриветмир世界こんにちは🙂"))

    @settings(max_examples=300, deadline=None)
    @given(st.text(alphabet=alphabet, max_size=60))
    def check(text):
        assert vocab.tokenize(text.encode("utf-8"), add_bos=False, special=False) == tok.encode(text, add_special_tokens=False).ids

    check()
