"""Paged KV-cache allocator (host side).

llama.cpp keeps one contiguous f16 K and V buffer per layer sized n_ctx and a slot finder
(`llama_kv_cache`, un-vendored upstream behind /root/reference/api.py:24-28).
This engine pages it: the device holds one pool of fixed-size pages shared by all sequences,
and each sequence owns a page table. This class hands out page ids; the engine only ever sees
`b2_seq_set_blocks(slot, pages)`.

Invariants (property-tested in tests/test_kv_pages.py): a page is owned by at most one sequence,
pages are never leaked, `free + owned == num_blocks`.
"""
from __future__ import annotations

import threading
from typing import Dict, List


class OutOfPages(RuntimeError):
    pass


class PageAllocator:
    def __init__(self, num_blocks: int, block_tokens: int):
        if num_blocks <= 0 or block_tokens <= 0:
            raise ValueError("num_blocks and block_tokens must be positive")
        self.num_blocks = num_blocks
        self.block_tokens = block_tokens
        self._free: List[int] = list(range(num_blocks - 1, -1, -1))   # pop() hands out low ids first
        self._owned: Dict[int, List[int]] = {}
        self._lock = threading.Lock()

    def pages_for(self, n_tokens: int) -> int:
        return (n_tokens + self.block_tokens - 1) // self.block_tokens

    @property
    def free_pages(self) -> int:
        return len(self._free)

    def table(self, seq_id: int) -> List[int]:
        return list(self._owned.get(seq_id, ()))

    def ensure(self, seq_id: int, n_tokens: int) -> bool:
        """Grow seq_id's page table to cover n_tokens. Returns True if the table changed."""
        with self._lock:
            pages = self._owned.setdefault(seq_id, [])
            need = self.pages_for(n_tokens) - len(pages)
            if need <= 0:
                return False
            if need > len(self._free):
                raise OutOfPages(f"need {need} KV pages, {len(self._free)} free")
            for _ in range(need):
                pages.append(self._free.pop())
            return True

    def release(self, seq_id: int) -> None:
        with self._lock:
            pages = self._owned.pop(seq_id, [])
            self._free.extend(reversed(pages))

    def truncate(self, seq_id: int, n_tokens: int) -> None:
        """Keep only the pages covering the first n_tokens (prefix reuse)."""
        with self._lock:
            pages = self._owned.get(seq_id, [])
            keep = self.pages_for(n_tokens)
            while len(pages) > keep:
                self._free.append(pages.pop())
