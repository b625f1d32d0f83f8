"""Python handle on one native engine (one GPU): GGUF -> HBM upload and the step calls.

Host-side half of what `Llama.__init__` does in the reference stack
(/root/reference/api.py:24-28 -> llama-cpp-python `Llama.__init__` -> llama.cpp model loader):
parse the file, read the `llama.*` hyper-parameters, copy every tensor's raw quant blocks
unchanged into device memory through the C ABI, size the paged KV pool.
"""
from __future__ import annotations

import ctypes
import os
import threading
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from . import _cabi
from ._cabi import B2Config, B2Sampling, check
from .gguf_file import GGUFFile
from .kv_pages import PageAllocator


@dataclass
class ModelDims:
    n_layer: int
    n_embd: int
    n_head: int
    n_head_kv: int
    head_dim: int
    n_ff: int
    n_vocab: int
    rms_eps: float
    rope_base: float
    n_ctx_train: int


@dataclass
class SamplingSettings:
    temperature: float = 0.8
    top_p: float = 0.95
    min_p: float = 0.05
    top_k: int = 40
    repeat_penalty: float = 1.1
    frequency_penalty: float = 0.0
    presence_penalty: float = 0.0
    penalty_last_n: int = 64

    def to_c(self) -> B2Sampling:
        return B2Sampling(self.temperature, self.top_p, self.min_p, self.repeat_penalty,
                          self.frequency_penalty, self.presence_penalty, self.top_k, self.penalty_last_n)


def read_dims(g: GGUFFile) -> ModelDims:
    md = g.metadata
    arch = md.get("general.architecture", "llama")
    if arch != "llama":
        raise ValueError(f"unsupported architecture {arch!r}: this engine implements the llama graph")
    k = lambda s, d=None: md.get(f"{arch}.{s}", d)
    n_embd = int(k("embedding_length"))
    n_head = int(k("attention.head_count"))
    n_head_kv = int(k("attention.head_count_kv", n_head))
    head_dim = int(k("rope.dimension_count", n_embd // n_head))
    if head_dim != n_embd // n_head:
        raise ValueError("partial rotary dimensions are not supported")
    scaling = k("rope.scaling.type")
    if scaling not in (None, "none"):
        raise ValueError(f"rope scaling {scaling!r} is not supported")
    return ModelDims(
        n_layer=int(k("block_count")), n_embd=n_embd, n_head=n_head, n_head_kv=n_head_kv, head_dim=head_dim,
        n_ff=int(k("feed_forward_length")), n_vocab=int(g.tensors["token_embd.weight"].shape[1]),
        rms_eps=float(k("attention.layer_norm_rms_epsilon", 1e-5)), rope_base=float(k("rope.freq_base", 10000.0)),
        n_ctx_train=int(k("context_length", 2048)))


class Engine:
    """One model on one GPU. Calls are serialised by `self.lock` (the native side also locks)."""

    def __init__(self, gguf: GGUFFile, n_ctx: int, max_seqs: int = 1, device: int = 0,
                 kv_block_tokens: int = 64, flags: int = 0, verbose: bool = False, tp=None):
        """`tp` = (rank, world, nccl_id: 128 bytes from `Engine.tp_unique_id()` of rank 0): tensor-parallel shard of the model
        (one process per GPU; every rank constructs its engine at the same time -- the set-up is a collective)."""
        self.lib = _cabi.load_library()
        self.dims = read_dims(gguf)
        self.tp_rank, self.tp_world = (int(tp[0]), int(tp[1])) if tp else (0, 1)
        full_dims = self.dims
        if self.tp_world > 1:
            from . import tp_shard
            self.dims = tp_shard.local_dims(full_dims, self.tp_world)     # local heads / n_ff; n_embd and n_vocab stay global
        self.n_ctx = int(n_ctx)
        self.max_seqs = int(max_seqs)
        self.device = int(device)
        self.block_tokens = int(kv_block_tokens)
        per_seq = (self.n_ctx + self.block_tokens - 1) // self.block_tokens
        self.pages = PageAllocator(per_seq * self.max_seqs, self.block_tokens)
        d = self.dims
        env_flags = 0
        if os.environ.get("B2_NO_GRAPH"):
            env_flags |= _cabi.B2_FLAG_NO_GRAPH
        if os.environ.get("B2_NO_PDL"):
            env_flags |= _cabi.B2_FLAG_NO_PDL
        if os.environ.get("B2_ACT8"):
            env_flags |= _cabi.B2_FLAG_ACT8
        cfg = B2Config(d.n_layer, d.n_embd, d.n_head, d.n_head_kv, d.head_dim, d.n_ff, d.n_vocab,
                       d.rms_eps, d.rope_base, self.n_ctx, self.max_seqs, self.block_tokens,
                       self.pages.num_blocks, self.device, flags | env_flags)
        h = ctypes.c_void_p()
        check(self.lib.b2_engine_create(ctypes.byref(cfg), ctypes.byref(h)), "b2_engine_create")
        self._h = h
        self.lock = threading.RLock()
        try:
            if self.tp_world > 1:
                idb = (ctypes.c_char * 128).from_buffer_copy(bytes(tp[2]))
                check(self.lib.b2_engine_tp_init(self._h, self.tp_rank, self.tp_world, idb), "b2_engine_tp_init")
                from . import tp_shard
                items = tp_shard.shard_all(gguf.tensors, self.tp_rank, self.tp_world)
            else:
                items = {name: (t.data, t.ggml_type, t.shape) for name, t in gguf.tensors.items()}
            for name, (data, ggml_type, shape) in items.items():
                ne0 = shape[0]
                ne1 = 1
                for s in shape[1:]:
                    ne1 *= s
                data = np.ascontiguousarray(data)
                check(self.lib.b2_upload_tensor(self._h, name.encode(), ggml_type, ne0, ne1,
                                                data.ctypes.data_as(ctypes.c_void_p), data.nbytes),
                      f"upload {name}")
            check(self.lib.b2_engine_finalize(self._h), "b2_engine_finalize")
        except Exception:
            self.close()
            raise
        self._tok = (ctypes.c_int32 * max(1, self.max_seqs))()
        self._slots = (ctypes.c_int32 * max(1, self.max_seqs))()

    @staticmethod
    def tp_unique_id() -> bytes:
        """128-byte NCCL id (rank 0 makes it, the host distributes it to the other ranks)."""
        buf = (ctypes.c_char * 128)()
        check(_cabi.load_library().b2_tp_unique_id(buf), "b2_tp_unique_id")
        return bytes(buf)

    # ---------------------------------------------------------------- lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self.lib.b2_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---------------------------------------------------------------- sequences
    def seq_begin(self, slot: int, n_past: int, history: Sequence[int], sampling: SamplingSettings,
                  uniforms: Optional[np.ndarray] = None):
        hist = list(history)[-64:]
        ha = (ctypes.c_int32 * max(1, len(hist)))(*hist)
        sp = sampling.to_c()
        if uniforms is None:
            uniforms = np.zeros(1, dtype=np.float64)
        u = np.ascontiguousarray(uniforms[:8192], dtype=np.float64)
        check(self.lib.b2_seq_begin(self._h, slot, n_past, ha, len(hist), ctypes.byref(sp),
                                    u.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), len(u)), "b2_seq_begin")

    def ensure_pages(self, slot: int, n_tokens: int):
        """Make sure the slot's page table covers n_tokens positions (uploads it if it grew)."""
        if self.pages.ensure(slot, n_tokens):
            tbl = self.pages.table(slot)
            arr = (ctypes.c_int32 * len(tbl))(*tbl)
            check(self.lib.b2_seq_set_blocks(self._h, slot, arr, len(tbl)), "b2_seq_set_blocks")

    def release(self, slot: int):
        self.pages.release(slot)

    # ---------------------------------------------------------------- forward
    def prefill(self, slot: int, tokens: Sequence[int]):
        arr = (ctypes.c_int32 * len(tokens))(*tokens)
        check(self.lib.b2_prefill(self._h, slot, arr, len(tokens)), "b2_prefill")

    def sample(self, slot: int) -> int:
        out = ctypes.c_int32()
        check(self.lib.b2_sample(self._h, slot, ctypes.byref(out)), "b2_sample")
        return out.value

    def decode(self, slots: Sequence[int]) -> List[int]:
        n = len(slots)
        for i, s in enumerate(slots):
            self._slots[i] = s
        check(self.lib.b2_decode(self._h, self._slots, n, self._tok), "b2_decode")
        return [self._tok[i] for i in range(n)]

    def decode_n(self, slot: int, n: int) -> List[int]:
        """n back-to-back decode steps of one slot; tokens come back in one copy."""
        out = (ctypes.c_int32 * n)()
        check(self.lib.b2_decode_n(self._h, slot, n, out), "b2_decode_n")
        return list(out)

    def rewind(self, slot: int, n_keep: int, pending_token: int, history: Sequence[int]):
        hist = list(history)[-64:]
        ha = (ctypes.c_int32 * max(1, len(hist)))(*hist)
        check(self.lib.b2_seq_rewind(self._h, slot, n_keep, pending_token, ha, len(hist)), "b2_seq_rewind")

    def set_token(self, slot: int, token: int):
        check(self.lib.b2_seq_set_token(self._h, slot, token), "b2_seq_set_token")

    def forward_pending(self, slot: int):
        check(self.lib.b2_forward_pending(self._h, slot), "b2_forward_pending")

    def logits(self, slot: int) -> np.ndarray:
        out = np.empty(self.dims.n_vocab, dtype=np.float32)
        check(self.lib.b2_get_logits(self._h, slot, out.ctypes.data_as(ctypes.c_void_p)), "b2_get_logits")
        return out

    # ---------------------------------------------------------------- measurement
    def timer_start(self):
        check(self.lib.b2_timer_start(self._h), "b2_timer_start")

    def timer_stop(self) -> float:
        ms = ctypes.c_float()
        check(self.lib.b2_timer_stop(self._h, ctypes.byref(ms)), "b2_timer_stop")
        return ms.value

    def launch_count(self) -> int:
        n = ctypes.c_int64()
        check(self.lib.b2_kernel_launch_count(self._h, ctypes.byref(n)))
        return n.value

    def mode(self) -> dict:
        v = [ctypes.c_int() for _ in range(5)]
        check(self.lib.b2_engine_mode(self._h, *[ctypes.byref(x) for x in v]))
        t = [ctypes.c_int() for _ in range(3)]
        check(self.lib.b2_engine_tp_mode(self._h, *[ctypes.byref(x) for x in t]))
        out = dict(graph=bool(v[0].value), pdl=bool(v[1].value), act_planes=v[2].value,
                   n_ctas=v[3].value, attn_splits=v[4].value)
        if t[1].value > 1:
            out.update(tp_rank=t[0].value, tp_world=t[1].value, tp_allreduce="peer memory" if t[2].value else "nccl")
        return out

    def device_bytes(self):
        w, kv = ctypes.c_int64(), ctypes.c_int64()
        check(self.lib.b2_engine_device_bytes(self._h, ctypes.byref(w), ctypes.byref(kv)))
        return w.value, kv.value

    def profile_decode_step(self, slot: int):
        ms = (ctypes.c_float * 8)()
        nl = (ctypes.c_int32 * 8)()
        by = (ctypes.c_int64 * 8)()
        check(self.lib.b2_profile_decode_step(self._h, slot, ms, nl, by), "b2_profile_decode_step")
        kinds = ["embed", "qkv", "attention", "attn_out", "gate_up", "ffn_down", "lm_head", "sampler"]
        return {k: dict(ms=ms[i], launches=nl[i], bytes=by[i]) for i, k in enumerate(kinds)}
