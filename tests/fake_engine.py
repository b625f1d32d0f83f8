"""CPU stand-in for llama_cpp.engine.Engine used by the host-logic tests (scheduler, Llama API,
reference api.py boundary). Deterministic: the next token is a hash of the tokens evaluated so
far, so batched and serial runs can be compared. No GPU, no oracle."""
import threading
import time

import numpy as np

from llama_cpp.engine import ModelDims, SamplingSettings
from llama_cpp.kv_pages import PageAllocator


class FakeEngine:
    delay = 0.0          # seconds per decode call (to exercise timeouts / back-pressure)
    eos_after = None     # emit `eos_token` after this many generated tokens
    eos_token = None
    fail_with = None     # raise this from prefill (error propagation tests)

    def __init__(self, gguf, n_ctx, max_seqs=1, device=0, kv_block_tokens=64, flags=0, verbose=False, tp=None):
        md = gguf.metadata
        n_embd, n_head = int(md["llama.embedding_length"]), int(md["llama.attention.head_count"])
        self.dims = ModelDims(int(md["llama.block_count"]), n_embd, n_head,
                              int(md.get("llama.attention.head_count_kv", n_head)), n_embd // n_head,
                              int(md["llama.feed_forward_length"]), int(gguf.tensors["token_embd.weight"].shape[1]),
                              1e-5, 10000.0, 2048)
        self.n_ctx, self.max_seqs, self.block_tokens = int(n_ctx), int(max_seqs), int(kv_block_tokens)
        per = (self.n_ctx + self.block_tokens - 1) // self.block_tokens
        self.pages = PageAllocator(per * self.max_seqs, self.block_tokens)
        self.lock = threading.RLock()
        self.seqs = {}
        self.calls = dict(prefill=0, prefill_tokens=0, decode=0, decode_n=0, rewind=0, max_batch=0, in_flight=0, max_in_flight=0)

    # -- helpers
    def _next(self, toks):
        h = 1469598103934665603
        for t in toks[-8:]:
            h = ((h ^ (t + 1)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return int(h % 200) + 300                      # stay inside the synthetic vocab's merged-token range

    def _sample(self, s):
        st = self.seqs[s]
        gen = len(st["toks"]) - st["n_prompt"]
        if self.eos_after is not None and gen >= self.eos_after:
            return self.eos_token
        return self._next(st["toks"])

    # -- Engine interface
    def ensure_pages(self, slot, n):
        self.pages.ensure(slot, n)

    def release(self, slot):
        self.pages.release(slot)

    def seq_begin(self, slot, n_past, history, sampling, uniforms=None):
        prev = self.seqs.get(slot, {"toks": []})["toks"]
        self.seqs[slot] = {"toks": list(prev[:n_past]), "pending": None, "n_prompt": 0}
        assert len(self.seqs[slot]["toks"]) == n_past, "prefix reuse asked for tokens that are not in the cache"

    def prefill(self, slot, tokens):
        if self.fail_with is not None:
            raise self.fail_with
        assert len(self.pages.table(slot)) * self.block_tokens >= len(self.seqs[slot]["toks"]) + len(tokens)
        self.calls["prefill"] += 1
        self.calls["prefill_tokens"] += len(tokens)
        st = self.seqs[slot]
        st["toks"].extend(int(t) for t in tokens)
        st["n_prompt"] = len(st["toks"])

    def sample(self, slot):
        t = self._sample(slot)
        self.seqs[slot]["pending"] = t
        return t

    def decode(self, slots):
        self.calls["decode"] += 1
        self.calls["max_batch"] = max(self.calls["max_batch"], len(slots))
        if self.delay:
            time.sleep(self.delay)
        out = []
        for s in slots:
            st = self.seqs[s]
            st["toks"].append(st["pending"])
            assert len(st["toks"]) <= self.n_ctx
            st["pending"] = self._sample(s)
            out.append(st["pending"])
        return out

    def decode_n(self, slot, n):
        self.calls["decode_n"] += 1
        return [self.decode([slot])[0] for _ in range(n)]

    def rewind(self, slot, n_keep, pending, history):
        self.calls["rewind"] += 1
        st = self.seqs[slot]
        st["toks"] = st["toks"][:n_keep]
        st["pending"] = pending

    def set_token(self, slot, token):
        self.seqs[slot]["pending"] = int(token)

    def forward_pending(self, slot):
        st = self.seqs[slot]
        st["toks"].append(st["pending"])

    def logits(self, slot):
        out = np.zeros(self.dims.n_vocab, np.float32)
        out[self._next(self.seqs[slot]["toks"])] = 1.0
        return out

    def mode(self):
        return dict(graph=False, pdl=False, act_planes=0, n_ctas=0, attn_splits=0)

    def close(self):
        pass
