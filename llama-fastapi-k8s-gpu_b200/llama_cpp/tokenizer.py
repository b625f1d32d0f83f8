"""GGUF vocabularies: SentencePiece-style ("llama") and byte-level BPE ("gpt2") tokenizers.

Host-side replacement for llama.cpp's `llm_tokenizer_spm` / `llm_tokenizer_bpe`, special-token
partitioning and `llama_token_to_piece`, which the reference reaches through
`llm.create_chat_completion` -> `Llama.tokenize` / `Llama.detokenize`
(/root/reference/api.py:55-63; un-vendored upstream). Tokenisation is not on
the per-token GPU path, so it stays in Python.

SPM: text -> (optional leading space) -> ' ' replaced by U+2581 -> UTF-8 characters -> greedily
merge the adjacent pair whose concatenation is the highest-scoring vocabulary piece (ties:
leftmost) -> pieces not in the vocabulary fall back to <0xXX> byte tokens.
BPE: text -> pre-tokeniser regex (llama-bpe = the Llama-3 pattern) -> bytes mapped to printable
code points (GPT-2 table) -> repeatedly merge the adjacent pair with the lowest merge rank.
"""
from __future__ import annotations

import heapq
from typing import Dict, List, Optional, Sequence, Tuple

# token types (llama.cpp llama_token_type)
TT_NORMAL, TT_UNKNOWN, TT_CONTROL, TT_USER_DEFINED, TT_UNUSED, TT_BYTE = 1, 2, 3, 4, 5, 6

LLAMA3_PRE_REGEX = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|"
                    r"\s*[\r\n]+|\s+(?!\S)|\s+")
GPT2_PRE_REGEX = r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"


def _gpt2_byte_table() -> Tuple[List[str], Dict[str, int]]:
    bs = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    enc = [""] * 256
    for b, c in zip(bs, cs):
        enc[b] = chr(c)
    return enc, {c: b for b, c in enumerate(enc)}


_BYTE_ENC, _BYTE_DEC = _gpt2_byte_table()


class Vocab:
    def __init__(self, metadata: dict, n_vocab_fallback: int = 0):
        md = metadata
        self.model = md.get("tokenizer.ggml.model", None)
        toks = md.get("tokenizer.ggml.tokens")
        if toks is None:
            # no vocabulary in the file (weights-only synthetic models): ids only
            self.model = "none"
            toks = []
        self.tokens: List[bytes] = [t if isinstance(t, bytes) else str(t).encode() for t in toks]
        self.n_vocab = len(self.tokens) or n_vocab_fallback
        self.token_to_id: Dict[bytes, int] = {}
        for i, t in enumerate(self.tokens):
            self.token_to_id.setdefault(t, i)
        sc = md.get("tokenizer.ggml.scores")
        self.scores = [float(x) for x in sc] if sc is not None else [0.0] * len(self.tokens)
        tt = md.get("tokenizer.ggml.token_type")
        self.types = [int(x) for x in tt] if tt is not None else [TT_NORMAL] * len(self.tokens)
        self.bos_id = int(md.get("tokenizer.ggml.bos_token_id", 1 if self.model == "llama" else -1))
        self.eos_id = int(md.get("tokenizer.ggml.eos_token_id", 2 if self.model == "llama" else -1))
        self.eot_id = int(md.get("tokenizer.ggml.eot_token_id", -1))
        self.unk_id = int(md.get("tokenizer.ggml.unknown_token_id", 0))
        if self.eot_id < 0:
            for name in (b"<|eot_id|>", b"<|im_end|>", b"<|end|>", b"<end_of_turn>", b"<|endoftext|>"):
                if name in self.token_to_id:
                    self.eot_id = self.token_to_id[name]
                    break
        ab = md.get("tokenizer.ggml.add_bos_token")
        # llama.cpp of the pinned vintage adds BOS unless the file says `add_bos_token = false`, for SPM and BPE vocabularies alike
        self.add_bos = bool(ab) if ab is not None else (self.bos_id >= 0)
        self.add_eos = bool(md.get("tokenizer.ggml.add_eos_token", False))
        self.add_space_prefix = bool(md.get("tokenizer.ggml.add_space_prefix", True))
        self.pre = md.get("tokenizer.ggml.pre", "default")
        self.merge_rank: Dict[Tuple[str, str], int] = {}
        for r, m in enumerate(md.get("tokenizer.ggml.merges", []) or []):
            s = m.decode("utf-8") if isinstance(m, bytes) else str(m)
            a, _, b = s.partition(" ")
            self.merge_rank.setdefault((a, b), r)
        self.special: Dict[bytes, int] = {
            t: i for i, t in enumerate(self.tokens) if self.types[i] in (TT_CONTROL, TT_USER_DEFINED) and t}
        self._special_sorted = sorted(self.special, key=len, reverse=True)
        self.byte_tokens: Dict[int, int] = {}
        if self.model == "llama":
            for b in range(256):
                i = self.token_to_id.get(f"<0x{b:02X}>".encode())
                if i is not None:
                    self.byte_tokens[b] = i
        self._pre_re = None
        self._special_re = None
        self._bpe_cache: Dict[str, Tuple[int, ...]] = {}      # pre-token -> ids (chat prompts repeat their words)
        # llama-bpe vocabularies set ignore_merges: a pre-token that is itself in the vocabulary is one token, whatever the merges say
        self.ignore_merges = self.pre in ("llama-bpe", "llama3", "llama-v3")

    # ------------------------------------------------------------------ encode
    def is_eog(self, tok: int) -> bool:
        return tok >= 0 and tok in (self.eos_id, self.eot_id)

    def _split_special(self, text: bytes) -> List[Tuple[bool, bytes]]:
        """[(is_special, fragment)], longest special tokens first (llama.cpp tokenizer_st_partition). One compiled
        alternation instead of a pass over the text per special token: a Llama-3 vocabulary has hundreds of them and
        the per-token passes cost a second per chat prompt."""
        if not self._special_sorted:
            return [(False, text)]
        if self._special_re is None:
            import re
            self._special_re = re.compile(b"(" + b"|".join(re.escape(sp) for sp in self._special_sorted) + b")")
        parts = self._special_re.split(text)
        return [(i % 2 == 1, part) for i, part in enumerate(parts) if part]

    def tokenize(self, text: bytes, add_bos: bool = True, special: bool = False) -> List[int]:
        if self.model == "none":
            raise RuntimeError("this GGUF file has no tokenizer vocabulary; pass token ids")
        out: List[int] = []
        if add_bos and self.add_bos and self.bos_id >= 0:
            out.append(self.bos_id)
        frags = self._split_special(text) if special else [(False, text)]
        prev_special = True
        for is_sp, frag in frags:
            if is_sp:
                out.append(self.special[frag])
                prev_special = True
                continue
            if not frag:
                continue
            s = frag.decode("utf-8", errors="replace")
            if self.model == "llama":
                if self.add_space_prefix and prev_special:
                    s = " " + s
                out.extend(self._spm(s.replace(" ", "▁")))
            else:
                out.extend(self._bpe(s))
            prev_special = False
        if add_bos and self.add_eos and self.eos_id >= 0:
            out.append(self.eos_id)
        return out

    def _spm(self, text: str) -> List[int]:
        syms = list(text)
        n = len(syms)
        if n == 0:
            return []
        prev = list(range(-1, n - 1))
        nxt = list(range(1, n + 1))
        nxt[-1] = -1
        heap: List[Tuple[float, int, int, str]] = []

        def push(l: int, r: int):
            if l < 0 or r < 0:
                return
            piece = syms[l] + syms[r]
            i = self.token_to_id.get(piece.encode("utf-8"))
            if i is not None:
                heapq.heappush(heap, (-self.scores[i], l, r, piece))

        for i in range(n - 1):
            push(i, i + 1)
        while heap:
            _, l, r, piece = heapq.heappop(heap)
            if syms[l] is None or syms[r] is None or nxt[l] != r or syms[l] + syms[r] != piece:
                continue
            syms[l] = piece
            syms[r] = None
            nxt[l] = nxt[r]
            if nxt[r] >= 0:
                prev[nxt[r]] = l
            push(prev[l], l)
            push(l, nxt[l])
        out: List[int] = []
        i = 0
        while i >= 0:
            s = syms[i]
            if s is not None:
                tid = self.token_to_id.get(s.encode("utf-8"))
                if tid is not None:
                    out.append(tid)
                else:
                    for b in s.encode("utf-8"):
                        out.append(self.byte_tokens.get(b, self.unk_id))
            i = nxt[i]
        return out

    def _bpe(self, text: str) -> List[int]:
        if self._pre_re is None:
            import regex
            self._pre_re = regex.compile(LLAMA3_PRE_REGEX if self.pre in ("llama-bpe", "llama3", "llama-v3") else GPT2_PRE_REGEX)
        out: List[int] = []
        cache = self._bpe_cache
        for word in self._pre_re.findall(text):
            hit = cache.get(word)
            if hit is not None:
                out.extend(hit)
                continue
            n0 = len(out)
            syms = [_BYTE_ENC[b] for b in word.encode("utf-8")]
            if self.ignore_merges:
                whole = self.token_to_id.get("".join(syms).encode("utf-8"))
                if whole is not None:
                    syms = []
                    out.append(whole)
            while len(syms) > 1:
                best, bi = None, -1
                for i in range(len(syms) - 1):
                    r = self.merge_rank.get((syms[i], syms[i + 1]))
                    if r is not None and (best is None or r < best):
                        best, bi = r, i
                if best is None:
                    break
                syms[bi:bi + 2] = [syms[bi] + syms[bi + 1]]
            for s in syms:
                tid = self.token_to_id.get(s.encode("utf-8"))
                if tid is not None:
                    out.append(tid)
                else:
                    for ch in s:
                        t2 = self.token_to_id.get(ch.encode("utf-8"))
                        if t2 is not None:
                            out.append(t2)
            if len(cache) < 200000:
                cache[word] = tuple(out[n0:])
        return out

    # ------------------------------------------------------------------ decode
    def token_to_piece(self, tok: int, special: bool = False) -> bytes:
        if tok < 0 or tok >= len(self.tokens):
            return b""
        ty = self.types[tok]
        t = self.tokens[tok]
        if ty in (TT_CONTROL, TT_UNKNOWN, TT_UNUSED):
            return t if (special and ty == TT_CONTROL) else b""
        if self.model == "llama":
            if ty == TT_BYTE:
                return bytes([int(t[3:5], 16)])
            return t.decode("utf-8", errors="replace").replace("▁", " ").encode("utf-8")
        if ty == TT_USER_DEFINED:
            return t
        s = t.decode("utf-8", errors="replace")
        return bytes(_BYTE_DEC.get(ch, ord("?") if ord(ch) > 255 else ord(ch)) for ch in s)

    def detokenize(self, tokens: Sequence[int], special: bool = False) -> bytes:
        if self.model == "none":
            return b"".join(b"<%d>" % t for t in tokens)
        out = b"".join(self.token_to_piece(t, special) for t in tokens)
        if self.model == "llama" and len(tokens) > 0 and tokens[0] == self.bos_id and out[:1] == b" ":
            out = out[1:]
        return out
