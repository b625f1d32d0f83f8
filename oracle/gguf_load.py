"""ORACLE (test infrastructure, not product code) -- GGUF container parsing for the oracle.

The oracle must not lean on the product's reader (llama_cpp/gguf_file.py), so it walks the file with
ggml's own Python package instead: `gguf.GGUFReader` (gguf-py 0.19, gguf/gguf_reader.py:132-370 -- header,
key/value section, tensor infos, 32-byte aligned data section). Every tensor comes back as the raw block
bytes exactly as they sit in the file, plus its ggml type id and its shape in ggml order (ne0 = the
contiguous dimension first); oracle/ggml_blocks.py turns the bytes into numbers.

Independent artefact that pins this module: gguf-py itself (it IS the reference implementation of the
container format); tests/test_gguf_file.py additionally checks the product's reader against it.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Tuple

import numpy as np


@dataclass
class OracleTensor:
    data: np.ndarray            # uint8, the tensor's bytes in file order (a view of the memory map)
    ggml_type: int
    shape: Tuple[int, ...]      # ggml order: shape[0] = K (contiguous), shape[1] = rows


class OracleGGUF:
    def __init__(self, path: str):
        import gguf
        self._reader = gguf.GGUFReader(path)           # keeps the memory map alive
        self.metadata: Dict[str, object] = {}
        for name, field in self._reader.fields.items():
            if name.startswith("GGUF."):
                continue
            if name.startswith("tokenizer.ggml.") and name not in ("tokenizer.ggml.model", "tokenizer.ggml.pre"):
                continue                                # vocabulary arrays: the forward pass does not need them
            self.metadata[name] = field.contents()
        self.tensors: Dict[str, OracleTensor] = {}
        for t in self._reader.tensors:
            raw = np.ascontiguousarray(t.data).view(np.uint8).reshape(-1)
            self.tensors[t.name] = OracleTensor(raw, int(t.tensor_type), tuple(int(s) for s in t.shape))
