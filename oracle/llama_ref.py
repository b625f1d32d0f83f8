"""ORACLE (test infrastructure, not product code) -- llama-architecture forward pass.

CPU float32 restatement of the graph the reference runs inside
`llm.create_chat_completion(...)` (/root/reference/api.py:55-63 -> llama-cpp-python
0.2.77 `Llama.eval` -> llama.cpp `llm_build_llama`; un-vendored upstream, SURVEY.md
section 2.2 H6). PARITY UNPINNED BY THE REFERENCE (no llama_cpp in this image, no
tests or golden vectors in the reference repo). Pinned instead against an independent
implementation that is importable here: transformers' GGUF loader -> float32
`LlamaForCausalLM` (tests/golden/make_golden.py writes the committed fixture,
tests/test_oracle_model.py compares).

Per layer (x = residual stream, float32):
    h   = rmsnorm(x, attn_norm, eps)
    q,k,v = Wq h, Wk h, Wv h
    q,k = rope(q), rope(k)      "norm" pairing: (x[2i], x[2i+1]) rotated by pos * base^(-2i/d)
    K,V cache append as float16 (upstream keeps an f16 KV cache)
    a   = softmax(q K^T / sqrt(d_head) + causal) V     (GQA: n_head / n_head_kv query heads per KV head)
    x   = x + Wo a
    h   = rmsnorm(x, ffn_norm, eps)
    x   = x + Wdown( silu(Wgate h) * (Wup h) )
logits = Woutput rmsnorm(x, output_norm)

Only tests/, smoke() and bench.py's CPU-baseline legs import this.
"""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np

from . import ggml_blocks as gb


@dataclass
class RefConfig:
    n_layer: int
    n_embd: int
    n_head: int
    n_head_kv: int
    n_ff: int
    n_vocab: int
    rms_eps: float
    rope_base: float

    @property
    def head_dim(self):
        return self.n_embd // self.n_head


def rope_freqs(head_dim: int, base: float) -> np.ndarray:
    """freq[i] = base^(-2i/d) built the way ggml does: theta_scale = powf(base, -2/d)
    multiplied up in float32. The CUDA engine uploads exactly this table."""
    scale = np.float32(np.float32(base) ** (np.float32(-2.0) / np.float32(head_dim)))
    out = np.empty(head_dim // 2, dtype=np.float32)
    f = np.float32(1.0)
    for i in range(head_dim // 2):
        out[i] = f
        f = np.float32(f * scale)
    return out


def rmsnorm(x: np.ndarray, w: np.ndarray, eps: float) -> np.ndarray:
    x = x.astype(np.float32)
    ms = np.mean(x.astype(np.float64) ** 2, axis=-1, keepdims=True)
    return (x * (1.0 / np.sqrt(ms + eps)).astype(np.float32) * w).astype(np.float32)


def rope(x: np.ndarray, pos: np.ndarray, freqs: np.ndarray) -> np.ndarray:
    """x [T, H, D]; adjacent-pair ("norm") rotation."""
    T, H, D = x.shape
    ang = pos.astype(np.float32)[:, None] * freqs[None, :]          # float32 product, like the kernel
    c = np.cos(ang.astype(np.float64)).astype(np.float32)[:, None, :]
    s = np.sin(ang.astype(np.float64)).astype(np.float32)[:, None, :]
    x0, x1 = x[..., 0::2], x[..., 1::2]
    out = np.empty_like(x)
    out[..., 0::2] = x0 * c - x1 * s
    out[..., 1::2] = x0 * s + x1 * c
    return out


class RefModel:
    """Float32 weights dequantised with the block oracle + a growing KV cache."""

    def __init__(self, gguf_path: str):
        # container format: ggml's own gguf-py reader (oracle/gguf_load.py), not the product's;
        # every number comes from oracle/ggml_blocks.py
        from .gguf_load import OracleGGUF
        g = OracleGGUF(gguf_path)
        md = g.metadata
        self.cfg = RefConfig(
            n_layer=int(md["llama.block_count"]), n_embd=int(md["llama.embedding_length"]),
            n_head=int(md["llama.attention.head_count"]),
            n_head_kv=int(md.get("llama.attention.head_count_kv", md["llama.attention.head_count"])),
            n_ff=int(md["llama.feed_forward_length"]),
            n_vocab=int(g.tensors["token_embd.weight"].shape[1]),
            rms_eps=float(md.get("llama.attention.layer_norm_rms_epsilon", 1e-5)),
            rope_base=float(md.get("llama.rope.freq_base", 10000.0)))
        self.w: Dict[str, np.ndarray] = {}
        for name, t in g.tensors.items():
            self.w[name] = gb.dequantize_tensor(t.data, t.ggml_type, t.shape)
        if "output.weight" not in self.w:
            self.w["output.weight"] = self.w["token_embd.weight"]
        self.freqs = rope_freqs(self.cfg.head_dim, self.cfg.rope_base)
        self.reset()

    def reset(self):
        c = self.cfg
        self.k_cache = [np.zeros((0, c.n_head_kv, c.head_dim), np.float16) for _ in range(c.n_layer)]
        self.v_cache = [np.zeros((0, c.n_head_kv, c.head_dim), np.float16) for _ in range(c.n_layer)]
        self.n_past = 0

    def forward(self, tokens, last_only: bool = False) -> np.ndarray:
        """Append `tokens` to the sequence; return logits [T, V] (or [1, V])."""
        c = self.cfg
        tokens = np.asarray(tokens, dtype=np.int64)
        T = len(tokens)
        pos = np.arange(self.n_past, self.n_past + T)
        x = self.w["token_embd.weight"][tokens].astype(np.float32)
        G = c.n_head // c.n_head_kv
        scale = np.float32(1.0 / np.sqrt(np.float32(c.head_dim)))
        for l in range(c.n_layer):
            p = f"blk.{l}."
            h = rmsnorm(x, self.w[p + "attn_norm.weight"], c.rms_eps)
            q = (h @ self.w[p + "attn_q.weight"].T).reshape(T, c.n_head, c.head_dim)
            k = (h @ self.w[p + "attn_k.weight"].T).reshape(T, c.n_head_kv, c.head_dim)
            v = (h @ self.w[p + "attn_v.weight"].T).reshape(T, c.n_head_kv, c.head_dim)
            q = rope(q, pos, self.freqs)
            k = rope(k, pos, self.freqs)
            self.k_cache[l] = np.concatenate([self.k_cache[l], k.astype(np.float16)], axis=0)
            self.v_cache[l] = np.concatenate([self.v_cache[l], v.astype(np.float16)], axis=0)
            K = self.k_cache[l].astype(np.float32)      # [S, Hkv, D]
            V = self.v_cache[l].astype(np.float32)
            S = K.shape[0]
            qg = q.reshape(T, c.n_head_kv, G, c.head_dim)
            s = np.einsum("thgd,shd->thgs", qg, K) * scale
            mask = np.arange(S)[None, :] > pos[:, None]          # [T, S]
            s = np.where(mask[:, None, None, :], -np.inf, s)
            s = s - s.max(axis=-1, keepdims=True)
            e = np.exp(s)
            pr = e / e.sum(axis=-1, keepdims=True)
            a = np.einsum("thgs,shd->thgd", pr.astype(np.float32), V).reshape(T, c.n_embd)
            x = x + a @ self.w[p + "attn_output.weight"].T
            h = rmsnorm(x, self.w[p + "ffn_norm.weight"], c.rms_eps)
            g = h @ self.w[p + "ffn_gate.weight"].T
            u = h @ self.w[p + "ffn_up.weight"].T
            act = (g / (1.0 + np.exp(-g))) * u
            x = x + act.astype(np.float32) @ self.w[p + "ffn_down.weight"].T
        self.n_past += T
        if last_only:
            x = x[-1:]
        h = rmsnorm(x, self.w["output_norm.weight"], c.rms_eps)
        return (h @ self.w["output.weight"].T).astype(np.float32)

    def greedy(self, prompt, n_new: int):
        """Greedy continuation; returns (tokens, logits per generated step [n_new, V])."""
        self.reset()
        lg = self.forward(prompt, last_only=True)[0]
        out, logs = [], []
        for _ in range(n_new):
            logs.append(lg)
            t = int(np.argmax(lg))
            out.append(t)
            lg = self.forward([t], last_only=True)[0]
        return out, np.stack(logs)
