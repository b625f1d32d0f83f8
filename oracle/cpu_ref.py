"""ORACLE / CPU BASELINE (test infrastructure, not product code) -- Python handle on cpu_ref.c.

Builds oracle/cpu_ref.c for the machine it runs on (gcc -O3 -march=native -fopenmp) into
oracle/_build/ and drives it through ctypes. Used by bench.py's `cpu_baseline` leg and
`--impl reference`, and by tests/test_cpu_ref.py. PARITY UNPINNED BY THE REFERENCE (see cpu_ref.c).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
from ctypes import POINTER, c_float, c_int, c_int32, c_int64, c_void_p

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")


class CrTensor(ctypes.Structure):
    _fields_ = [("data", c_void_p), ("type", c_int32), ("pad", c_int32), ("ne0", c_int64), ("ne1", c_int64)]


class CrModel(ctypes.Structure):
    _fields_ = [(n, c_int32) for n in ("n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab", "n_ctx")] + [
        ("rms_eps", c_float), ("rope_base", c_float), ("tok_embd", CrTensor), ("out_norm", CrTensor),
        ("output", CrTensor), ("layers", POINTER(CrTensor))]


def build(force: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    src = os.path.join(HERE, "cpu_ref.c")
    out = os.path.join(BUILD, "libcpu_ref.so")
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        cmd = ["gcc", "-O3", "-march=native", "-fopenmp", "-shared", "-fPIC", src, "-o", out, "-lm"]
        subprocess.run(cmd, check=True)
    return out


class CpuRef:
    def __init__(self, gguf_path: str, n_ctx: int, n_threads: int = 0):
        from .gguf_load import OracleGGUF          # gguf-py reader, independent of the product's
        try:
            self.lib = ctypes.CDLL(build())
        except OSError:
            self.lib = ctypes.CDLL(build(force=True))
        self.lib.cr_create.restype = c_void_p
        self.lib.cr_create.argtypes = [POINTER(CrModel)]
        self.lib.cr_eval.restype = POINTER(c_float)
        self.lib.cr_eval.argtypes = [c_void_p, c_int, c_int]
        self.lib.cr_reset.argtypes = [c_void_p]
        self.lib.cr_destroy.argtypes = [c_void_p]
        self.lib.cr_threads.restype = c_int
        self.lib.cr_set_threads.argtypes = [c_int]
        self.lib.cr_set_simd.argtypes = [c_int]
        self.lib.cr_set_exact.argtypes = [c_int]
        self.lib.cr_simd.restype = c_int
        if n_threads:
            self.lib.cr_set_threads(int(n_threads))
        self.g = OracleGGUF(gguf_path)
        md = self.g.metadata

        def T(name):
            t = self.g.tensors[name]
            ne1 = 1
            for s in t.shape[1:]:
                ne1 *= s
            return CrTensor(t.data.ctypes.data, t.ggml_type, 0, t.shape[0], ne1)

        L = int(md["llama.block_count"])
        self._layers = (CrTensor * (9 * L))()
        names = ["attn_norm", "attn_q", "attn_k", "attn_v", "attn_output", "ffn_norm", "ffn_gate", "ffn_up", "ffn_down"]
        for l in range(L):
            for j, n in enumerate(names):
                self._layers[9 * l + j] = T(f"blk.{l}.{n}.weight")
        n_embd, n_head = int(md["llama.embedding_length"]), int(md["llama.attention.head_count"])
        self.n_vocab = int(self.g.tensors["token_embd.weight"].shape[1])
        out = "output.weight" if "output.weight" in self.g.tensors else "token_embd.weight"
        self.model = CrModel(L, n_embd, n_head, int(md.get("llama.attention.head_count_kv", n_head)), n_embd // n_head,
                             int(md["llama.feed_forward_length"]), self.n_vocab, n_ctx,
                             float(md.get("llama.attention.layer_norm_rms_epsilon", 1e-5)),
                             float(md.get("llama.rope.freq_base", 10000.0)), T("token_embd.weight"),
                             T("output_norm.weight"), T(out), self._layers)
        self.ctx = self.lib.cr_create(ctypes.byref(self.model))
        self.threads = self.lib.cr_threads()

    @property
    def simd(self) -> str:
        return "avx512-vnni" if self.lib.cr_simd() else "scalar"

    @property
    def has_vnni(self) -> bool:
        return bool(self.lib.cr_has_vnni())

    def set_simd(self, on: bool):
        self.lib.cr_set_simd(1 if on else 0)

    def set_exact(self, on: bool):
        """Checker mode: float activations against dequantised weights with double accumulation (no int8 activation noise).
        Process-wide, like set_simd."""
        self.lib.cr_set_exact(1 if on else 0)

    def reset(self):
        self.lib.cr_reset(self.ctx)

    def eval(self, token: int, want_logits: bool = True):
        p = self.lib.cr_eval(self.ctx, int(token), 1 if want_logits else 0)
        if not p:
            raise ValueError("cr_eval: bad token or context full")
        if not want_logits:
            return None
        return np.ctypeslib.as_array(p, shape=(self.n_vocab,)).copy()

    def close(self):
        if self.ctx:
            self.lib.cr_destroy(self.ctx)
            self.ctx = None
