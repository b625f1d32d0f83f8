"""Seeded synthetic GGUF models of the architectures BASELINE.json names.

There is no network and no checkpoint on the GPU box, so throughput and parity
runs use random-init GGUF files with the tensor-type mix of the real llama.cpp
ftypes: quant nibbles/bytes are uniform random, the
fp16 scale fields are overwritten so every tensor dequantises to roughly
zero-mean values of a chosen std (0.02 by default, like a fresh transformer).

The file is a normal GGUF: the same loader path (`gguf_file.GGUFFile` ->
C-ABI upload) serves it that would serve `Lexi-Llama-3-8B-Uncensored_Q4_K_M.gguf`
(/root/reference/api.py:13-15).
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np

from .gguf_file import GGML_TYPE_ID, GGML_TYPES, write_gguf


@dataclass
class Arch:
    name: str
    n_layer: int
    n_embd: int
    n_head: int
    n_head_kv: int
    n_ff: int
    n_vocab: int
    rope_base: float = 10000.0
    rms_eps: float = 1e-5
    n_ctx_train: int = 8192
    tokenizer: str = "bpe"  # "bpe" (llama-3 style byte-level) or "spm"

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head


ARCHS: Dict[str, Arch] = {
    # tiny test architectures (CPU oracle finishes in well under a second)
    "tiny": Arch("tiny", 2, 512, 8, 2, 1536, 1024, 10000.0, 1e-5, 2048, "bpe"),
    "tiny128": Arch("tiny128", 2, 512, 4, 2, 1024, 768, 500000.0, 1e-5, 2048, "bpe"),
    # one layer of the benchmark geometries: the launch shapes the 8B / 70B runs use (row counts per CTA, K slices,
    # the Q6_K attn_v / ffn_down / output of the Q4_K_M mix, a 128256-row output head) at a size the float32
    # oracle still handles (tests/test_gpu_wide.py)
    "wide-1l": Arch("wide-1l", 1, 4096, 32, 8, 14336, 128256, 500000.0, 1e-5, 8192, "bpe"),
    "wide70-1l": Arch("wide70-1l", 1, 8192, 64, 8, 28672, 16384, 500000.0, 1e-5, 8192, "bpe"),
    "tinyllama-1.1b": Arch("tinyllama-1.1b", 22, 2048, 32, 4, 5632, 32000, 10000.0, 1e-5, 2048, "spm"),
    "llama-3-8b": Arch("llama-3-8b", 32, 4096, 32, 8, 14336, 128256, 500000.0, 1e-5, 8192, "bpe"),
    "mistral-7b": Arch("mistral-7b", 32, 4096, 32, 8, 14336, 32000, 10000.0, 1e-5, 32768, "spm"),
    "llama-3-70b": Arch("llama-3-70b", 80, 8192, 64, 8, 28672, 128256, 500000.0, 1e-5, 8192, "bpe"),
}

# std of the dequantised values when the block's fp16 scale field(s) are 1.0 and
# every other byte is uniform random (measured once against gguf-py's dequantize):
#   Q4_0  d=1                    -> 4.611
#   Q8_0  d=1                    -> 73.91
#   Q4_K  d=1, dmin=7.5          -> 258.6   (dmin = 7.5 d centres the values)
#   Q5_K  d=1, dmin=15.5         -> 526.7
#   Q6_K  d=1, scales in [-32,32]-> 347.0
#   Q4_1  d=1, m=-7.5            -> 4.612
#   Q5_0  d=1                    -> 9.234
#   Q5_1  d=1, m=-15.5           -> 9.228
#   Q2_K  d=1, dmin=1.5          -> 13.87
#   Q3_K  d=1                    -> 43.34
_UNIT_STD = {"Q4_0": 4.611, "Q8_0": 73.91, "Q4_K": 258.6, "Q5_K": 526.7, "Q6_K": 347.0,
             "Q4_1": 4.612, "Q5_0": 9.234, "Q5_1": 9.228, "Q2_K": 13.87, "Q3_K": 43.34}


def _random_bytes(n: int, seed: int) -> np.ndarray:
    """n uniform random bytes, generated on all host cores (about 6 GB/s on 8 cores)."""
    n8 = (n + 7) // 8
    out = np.empty(n8, dtype=np.uint64)
    chunk = 1 << 23  # 64 MiB of bytes per job
    jobs = [(i, min(chunk, n8 - i)) for i in range(0, n8, chunk)]

    def fill(job):
        i, m = job
        bg = np.random.SFC64([seed, i])
        out[i:i + m] = bg.random_raw(m)

    if len(jobs) == 1:
        fill(jobs[0])
    else:
        with ThreadPoolExecutor(min(8, os.cpu_count() or 1)) as ex:
            list(ex.map(fill, jobs))
    return out.view(np.uint8)[:n]


def _f16_bytes(vals: np.ndarray) -> np.ndarray:
    return vals.astype(np.float16).view(np.uint8).reshape(vals.shape + (2,))


def random_blocks(type_name: str, n_blocks: int, std: float, seed: int) -> np.ndarray:
    """(n_blocks, block_bytes) uint8 of random quant blocks dequantising to ~N(0, std)."""
    _, be, bb = GGML_TYPES[GGML_TYPE_ID[type_name]]
    if type_name == "F32":
        r = np.random.default_rng(seed).standard_normal(n_blocks).astype(np.float32) * std
        return r.view(np.uint8).reshape(n_blocks, 4)
    if type_name == "F16":
        r = np.random.default_rng(seed).standard_normal(n_blocks).astype(np.float16) * np.float16(std)
        return r.view(np.uint8).reshape(n_blocks, 2)
    b = _random_bytes(n_blocks * bb, seed).reshape(n_blocks, bb)
    rng = np.random.default_rng(seed ^ 0x5EED)
    d = (std / _UNIT_STD[type_name]) * rng.uniform(0.5, 1.5, size=n_blocks)
    if type_name in ("Q4_0", "Q8_0", "Q5_0"):
        d *= rng.choice([-1.0, 1.0], size=n_blocks)
        b[:, 0:2] = _f16_bytes(d)
    elif type_name in ("Q4_1", "Q5_1"):
        b[:, 0:2] = _f16_bytes(d)
        b[:, 2:4] = _f16_bytes((-7.5 if type_name == "Q4_1" else -15.5) * d)
    elif type_name == "Q2_K":
        b[:, 80:82] = _f16_bytes(d)
        b[:, 82:84] = _f16_bytes(1.5 * d)
    elif type_name == "Q3_K":
        b[:, 108:110] = _f16_bytes(d)
    elif type_name == "Q4_K":
        b[:, 0:2] = _f16_bytes(d)
        b[:, 2:4] = _f16_bytes(7.5 * d)
    elif type_name == "Q5_K":
        b[:, 0:2] = _f16_bytes(d)
        b[:, 2:4] = _f16_bytes(15.5 * d)
    elif type_name == "Q6_K":
        b[:, 192:208] = rng.integers(-32, 33, size=(n_blocks, 16)).astype(np.int8).view(np.uint8)
        b[:, 208:210] = _f16_bytes(d)
    else:
        raise ValueError(type_name)
    return b


def _use_more_bits(i: int, n: int) -> bool:
    """llama.cpp's Q4_K_M rule for which layers keep attn_v / ffn_down at Q6_K
    (restated from SURVEY.md section 8(d): i < n/8 or i >= 7n/8 or (i - n/8) % 3 == 2)."""
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_types(arch: Arch, ftype: str) -> Dict[str, str]:
    """tensor name -> ggml type name for a whole model of quantisation mix `ftype`."""
    L = arch.n_layer
    out: Dict[str, str] = {}

    def put(name, t):
        out[name] = t

    if ftype == "Q4_0":
        base, outw = "Q4_0", "Q6_K"
    elif ftype == "Q8_0":
        base, outw = "Q8_0", "Q8_0"
    elif ftype == "Q4_K_M":
        base, outw = "Q4_K", "Q6_K"
    elif ftype == "Q4_K":   # pure, for tests
        base, outw = "Q4_K", "Q4_K"
    elif ftype == "Q6_K":
        base, outw = "Q6_K", "Q6_K"
    elif ftype in ("Q5_K", "Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K"):   # pure, for tests
        base, outw = ftype, ftype
    elif ftype == "Q2_K_mix":      # like llama.cpp's Q2_K files: Q2_K with attn_v at Q4_K, ffn_down at Q3_K, output at Q6_K
        base, outw = "Q2_K", "Q6_K"
    elif ftype == "F16":
        base, outw = "F16", "F16"
    else:
        raise ValueError(f"unknown ftype {ftype}")
    put("token_embd.weight", base)
    put("output_norm.weight", "F32")
    put("output.weight", outw)
    for i in range(L):
        p = f"blk.{i}."
        put(p + "attn_norm.weight", "F32")
        put(p + "ffn_norm.weight", "F32")
        for t in ("attn_q", "attn_k", "attn_output", "ffn_gate", "ffn_up"):
            put(p + t + ".weight", base)
        more = ftype == "Q4_K_M" and _use_more_bits(i, L)
        v_t = "Q6_K" if more else base
        if ftype == "Q4_K_M" and not more and arch.n_head // arch.n_head_kv >= 8 and arch.n_layer >= 80:
            v_t = "Q5_K"  # 70B-class rule
        put(p + "attn_v.weight", "Q4_K" if ftype == "Q2_K_mix" else v_t)
        put(p + "ffn_down.weight", "Q3_K" if ftype == "Q2_K_mix" else "Q6_K" if more else base)
    return out


def tensor_shapes(arch: Arch) -> Dict[str, Tuple[int, ...]]:
    """ggml-order shapes: (K contiguous, N rows)."""
    d, ff, V = arch.n_embd, arch.n_ff, arch.n_vocab
    kv = arch.n_head_kv * arch.head_dim
    s: Dict[str, Tuple[int, ...]] = {
        "token_embd.weight": (d, V), "output_norm.weight": (d,), "output.weight": (d, V)}
    for i in range(arch.n_layer):
        p = f"blk.{i}."
        s[p + "attn_norm.weight"] = (d,)
        s[p + "ffn_norm.weight"] = (d,)
        s[p + "attn_q.weight"] = (d, d)
        s[p + "attn_k.weight"] = (d, kv)
        s[p + "attn_v.weight"] = (d, kv)
        s[p + "attn_output.weight"] = (d, d)
        s[p + "ffn_gate.weight"] = (d, ff)
        s[p + "ffn_up.weight"] = (d, ff)
        s[p + "ffn_down.weight"] = (ff, d)
    return s


# ---------------------------------------------------------------------------
# synthetic vocabularies
# ---------------------------------------------------------------------------
def gpt2_byte_chars() -> List[str]:
    """GPT-2 byte<->unicode table: printable stand-in character for each byte."""
    bs = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    table = dict(zip(bs, cs))
    return [chr(table[b]) for b in range(256)]


LLAMA3_SPECIALS = {
    128000: "<|begin_of_text|>", 128001: "<|end_of_text|>", 128006: "<|start_header_id|>",
    128007: "<|end_header_id|>", 128009: "<|eot_id|>",
}
LLAMA3_CHAT_TEMPLATE = (
    "{% set loop_messages = messages %}{% for message in loop_messages %}{% set content = "
    "'<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n'+ message['content'] | trim + "
    "'<|eot_id|>' %}{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}"
    "{{ content }}{% endfor %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}")


def synth_bpe_vocab(n_vocab: int):
    """Byte-level BPE vocab: 256 byte tokens, a few hundred merges of common English
    bigrams/words, filler user-defined tokens, and Llama-3's special tokens at their
    real ids when the vocab is large enough (else at the tail)."""
    bc = gpt2_byte_chars()
    enc = lambda s: "".join(bc[b] for b in s.encode("utf-8"))
    tokens: List[str] = list(bc)
    merges: List[str] = []
    have = set(tokens)
    words = ("the of and to in is you that it he was for on are as with his they at be this have "
             "from or one had by word but not what all were we when your can said there use an each "
             "which she do how their if will up other about out many then them these so some her "
             "would make like him into time has look two more write go see number no way could people "
             "my than first water been call who oil its now find long down day did get come made may "
             "part hello world user assistant system").split()
    budget = max(0, min(n_vocab - 256 - 16, 4000))
    for w in words:
        for form in (" " + w, w):
            e = enc(form)
            # left-to-right chain of merges building `e`
            parts = list(e)
            cur = parts[0]
            for ch in parts[1:]:
                nxt = cur + ch
                if nxt not in have:
                    if len(tokens) - 256 >= budget:
                        break
                    tokens.append(nxt)
                    have.add(nxt)
                    merges.append(f"{cur} {ch}")
                cur = nxt
    n_special = 5
    types = [1] * len(tokens)  # NORMAL
    specials = dict(LLAMA3_SPECIALS) if n_vocab >= 128010 else {
        n_vocab - 5 + i: s for i, s in enumerate(LLAMA3_SPECIALS.values())}
    while len(tokens) < n_vocab:
        i = len(tokens)
        if i in specials:
            tokens.append(specials[i])
            types.append(3)  # CONTROL
        else:
            tokens.append(f"<|reserved_{i}|>")
            types.append(1)  # NORMAL filler (renders as its text, never produced by BPE of plain text: no merge builds it).
            # Not USER_DEFINED: special-token parsing scans for every CONTROL / USER_DEFINED entry, and 127k of them is
            # nothing like a real Llama-3 vocabulary (256 control tokens)
    inv = {s: i for i, s in specials.items()}
    meta = {
        "tokenizer.ggml.model": "gpt2",
        "tokenizer.ggml.pre": "llama-bpe",
        "tokenizer.ggml.tokens": [t.encode("utf-8") for t in tokens],
        "tokenizer.ggml.token_type": np.asarray(types, dtype=np.int32),
        "tokenizer.ggml.merges": [m.encode("utf-8") for m in merges],
        "tokenizer.ggml.bos_token_id": inv["<|begin_of_text|>"],
        "tokenizer.ggml.eos_token_id": inv["<|end_of_text|>"],
        "tokenizer.ggml.eot_token_id": inv["<|eot_id|>"],
        "tokenizer.chat_template": LLAMA3_CHAT_TEMPLATE,
    }
    assert n_special == len(specials)
    return meta


def synth_spm_vocab(n_vocab: int):
    """SentencePiece-style vocab. Uses the real 32000-piece model that ships inside
    `mistral_common` when it is importable (offline), else a synthetic one with the
    same structure: <unk>, <s>, </s>, 256 byte pieces, then scored pieces."""
    pieces: List[Tuple[str, float, int]] = []
    try:
        import mistral_common  # type: ignore
        from sentencepiece import sentencepiece_model_pb2 as pb  # type: ignore
        path = os.path.join(os.path.dirname(mistral_common.__file__), "data", "tokenizer.model.v1")
        m = pb.ModelProto()
        with open(path, "rb") as f:
            m.ParseFromString(f.read())
        for p in m.pieces:
            pieces.append((p.piece, float(p.score), int(p.type)))
    except Exception:
        pieces = []
    if len(pieces) < 259:
        pieces = [("<unk>", 0.0, 2), ("<s>", 0.0, 3), ("</s>", 0.0, 3)]
        pieces += [(f"<0x{b:02X}>", 0.0, 6) for b in range(256)]
        base = "▁ e t a o i n s h r d l u ▁t ▁a he in ▁the re on er ▁s at ▁w ▁o en it ou ▁c is an or ▁b ▁hello ▁world".split()
        for k, p in enumerate(base):
            pieces.append((p, -float(k + 1), 1))
    pieces = pieces[:n_vocab]
    k = 0
    while len(pieces) < n_vocab:
        pieces.append((f"<extra_{k}>", -1e4 - k, 4))
        k += 1
    meta = {
        "tokenizer.ggml.model": "llama",
        "tokenizer.ggml.tokens": [p[0].encode("utf-8") for p in pieces],
        "tokenizer.ggml.scores": np.asarray([p[1] for p in pieces], dtype=np.float32),
        "tokenizer.ggml.token_type": np.asarray([p[2] for p in pieces], dtype=np.int32),
        "tokenizer.ggml.bos_token_id": 1,
        "tokenizer.ggml.eos_token_id": 2,
        "tokenizer.ggml.unknown_token_id": 0,
        "tokenizer.ggml.add_bos_token": True,
        "tokenizer.ggml.add_eos_token": False,
    }
    return meta


# ---------------------------------------------------------------------------
# model file
# ---------------------------------------------------------------------------
def write_synthetic_model(path: str, arch_name: str, ftype: str, seed: int = 0,
                          std: float = 0.02, output_std: Optional[float] = None,
                          n_ctx_train: Optional[int] = None, with_vocab: bool = True) -> str:
    """Write a random-init GGUF of architecture `arch_name` and quant mix `ftype`."""
    arch = ARCHS[arch_name]
    types = tensor_types(arch, ftype)
    shapes = tensor_shapes(arch)
    meta = {
        "general.architecture": "llama",
        "general.name": f"synthetic-{arch.name}-{ftype}-seed{seed}",
        "general.alignment": 32,
        "llama.block_count": arch.n_layer,
        "llama.context_length": int(n_ctx_train or arch.n_ctx_train),
        "llama.embedding_length": arch.n_embd,
        "llama.feed_forward_length": arch.n_ff,
        "llama.attention.head_count": arch.n_head,
        "llama.attention.head_count_kv": arch.n_head_kv,
        "llama.attention.layer_norm_rms_epsilon": float(arch.rms_eps),
        "llama.rope.freq_base": float(arch.rope_base),
        "llama.rope.dimension_count": arch.head_dim,
        "llama.vocab_size": arch.n_vocab,
    }
    if with_vocab:
        meta.update(synth_bpe_vocab(arch.n_vocab) if arch.tokenizer == "bpe"
                    else synth_spm_vocab(arch.n_vocab))
    tensors = []
    for idx, (name, shape) in enumerate(shapes.items()):
        tname = types[name]
        tid = GGML_TYPE_ID[tname]
        n = 1
        for d in shape:
            n *= d
        be = GGML_TYPES[tid][1]

        def make(tname=tname, n=n, be=be, name=name, idx=idx):
            if name.endswith("_norm.weight"):
                r = np.random.default_rng([seed, idx]).uniform(0.5, 1.5, size=n).astype(np.float32)
                return r.view(np.uint8)
            s = std
            if name == "output.weight" and output_std is not None:
                s = output_std
            return random_blocks(tname, n // be, s, seed * 100003 + idx).reshape(-1)

        tensors.append((name, tid, shape, make))
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    write_gguf(tmp, meta, tensors)
    os.replace(tmp, path)
    return path
