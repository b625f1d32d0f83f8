"""mmap GGUF (v2/v3) reader and a minimal writer.

Replaces the loader the reference reaches through ``Llama(model_path=...)``
(/root/reference/api.py:24-28 -> llama-cpp-python 0.2.77 -> llama.cpp
``llama_model_loader`` / ggml ``gguf_init_from_file``; un-vendored upstream).

File layout (little endian) as published by ggml and implemented in gguf-py
(``gguf/gguf_reader.py:132-185`` header/KV/tensor-info walk,
``gguf/constants.py:10-12`` magic/version/alignment):

    u32 magic 'GGUF' | u32 version | u64 n_tensors | u64 n_kv
    n_kv   x { string key | u32 vtype | value }
    n_tens x { string name | u32 n_dims | u64 dims[n_dims] | u32 ggml_type | u64 offset }
    pad to ``general.alignment`` (default 32)
    tensor data; each tensor at data_start + offset

Tensor bytes are never copied or re-packed here: ``GGUFTensor.data`` is a
zero-copy ``memoryview``/numpy view into the mmap, handed unchanged to the
C-ABI upload call.
"""
from __future__ import annotations

import mmap
import struct
from dataclasses import dataclass
from typing import Any, Dict, List, Sequence, Tuple

import numpy as np

GGUF_MAGIC = 0x46554747  # 'GGUF'
GGUF_DEFAULT_ALIGNMENT = 32

# ggml tensor types this engine understands: id -> (name, block elems, block bytes)
GGML_TYPES: Dict[int, Tuple[str, int, int]] = {
    0: ("F32", 1, 4),
    1: ("F16", 1, 2),
    2: ("Q4_0", 32, 18),
    3: ("Q4_1", 32, 20),
    6: ("Q5_0", 32, 22),
    7: ("Q5_1", 32, 24),
    8: ("Q8_0", 32, 34),
    10: ("Q2_K", 256, 84),
    11: ("Q3_K", 256, 110),
    12: ("Q4_K", 256, 144),
    13: ("Q5_K", 256, 176),
    14: ("Q6_K", 256, 210),
}
GGML_TYPE_ID = {v[0]: k for k, v in GGML_TYPES.items()}

# GGUF metadata value types
(_U8, _I8, _U16, _I16, _U32, _I32, _F32, _BOOL, _STR, _ARR, _U64, _I64, _F64) = range(13)
_SCALAR_FMT = {
    _U8: "<B", _I8: "<b", _U16: "<H", _I16: "<h", _U32: "<I", _I32: "<i",
    _F32: "<f", _BOOL: "<?", _U64: "<Q", _I64: "<q", _F64: "<d",
}
_NP_FMT = {
    _U8: np.uint8, _I8: np.int8, _U16: np.uint16, _I16: np.int16, _U32: np.uint32,
    _I32: np.int32, _F32: np.float32, _BOOL: np.bool_, _U64: np.uint64,
    _I64: np.int64, _F64: np.float64,
}


def tensor_nbytes(ggml_type: int, n_elems: int) -> int:
    if ggml_type not in GGML_TYPES:
        raise ValueError(f"unsupported ggml tensor type id {ggml_type}")
    _, be, bb = GGML_TYPES[ggml_type]
    if n_elems % be:
        raise ValueError(f"{n_elems} elements is not a whole number of {be}-element blocks")
    return n_elems // be * bb


@dataclass
class GGUFTensor:
    name: str
    ggml_type: int
    shape: Tuple[int, ...]  # ggml order: shape[0] = contiguous (K), shape[1] = rows (N)
    offset: int             # absolute file offset
    nbytes: int
    data: np.ndarray        # uint8 view into the mmap (zero copy)

    @property
    def type_name(self) -> str:
        return GGML_TYPES[self.ggml_type][0]

    @property
    def n_elems(self) -> int:
        n = 1
        for d in self.shape:
            n *= d
        return n


class GGUFFile:
    """Parsed GGUF file. ``metadata`` maps key -> python value (arrays -> list/np)."""

    def __init__(self, path: str):
        self.path = path
        self._f = open(path, "rb")
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        self._buf = memoryview(self._mm)
        self.metadata: Dict[str, Any] = {}
        self.tensors: Dict[str, GGUFTensor] = {}
        self._parse()

    # -- low level -----------------------------------------------------
    def _rd(self, fmt: str, off: int):
        return struct.unpack_from(fmt, self._buf, off)[0], off + struct.calcsize(fmt)

    def _rd_str(self, off: int):
        n, off = self._rd("<Q", off)
        s = bytes(self._buf[off:off + n])
        return s, off + n

    def _rd_value(self, vtype: int, off: int):
        if vtype in _SCALAR_FMT:
            return self._rd(_SCALAR_FMT[vtype], off)
        if vtype == _STR:
            s, off = self._rd_str(off)
            return s.decode("utf-8", errors="replace"), off
        if vtype == _ARR:
            itype, off = self._rd("<I", off)
            count, off = self._rd("<Q", off)
            if itype in _NP_FMT:
                dt = np.dtype(_NP_FMT[itype]).newbyteorder("<")
                arr = np.frombuffer(self._buf, dtype=dt, count=count, offset=off)
                return arr, off + count * dt.itemsize
            if itype == _STR:
                out: List[bytes] = []
                for _ in range(count):
                    s, off = self._rd_str(off)
                    out.append(s)
                return out, off  # token/merge strings kept as raw bytes
            if itype == _ARR:
                out2 = []
                for _ in range(count):
                    v, off = self._rd_value(_ARR, off)
                    out2.append(v)
                return out2, off
            raise ValueError(f"bad GGUF array item type {itype}")
        raise ValueError(f"bad GGUF value type {vtype}")

    def _parse(self):
        magic, off = self._rd("<I", 0)
        if magic != GGUF_MAGIC:
            raise ValueError(f"{self.path}: not a GGUF file (magic {magic:#x})")
        self.version, off = self._rd("<I", off)
        if self.version not in (2, 3):
            raise ValueError(f"{self.path}: unsupported GGUF version {self.version}")
        n_tensors, off = self._rd("<Q", off)
        n_kv, off = self._rd("<Q", off)
        for _ in range(n_kv):
            key, off = self._rd_str(off)
            vtype, off = self._rd("<I", off)
            val, off = self._rd_value(vtype, off)
            self.metadata[key.decode("utf-8")] = val
        infos = []
        for _ in range(n_tensors):
            name, off = self._rd_str(off)
            n_dims, off = self._rd("<I", off)
            dims = []
            for _ in range(n_dims):
                d, off = self._rd("<Q", off)
                dims.append(int(d))
            ttype, off = self._rd("<I", off)
            toff, off = self._rd("<Q", off)
            infos.append((name.decode("utf-8"), tuple(dims), int(ttype), int(toff)))
        align = int(self.metadata.get("general.alignment", GGUF_DEFAULT_ALIGNMENT))
        self.alignment = align
        self.data_start = (off + align - 1) // align * align
        whole = np.frombuffer(self._buf, dtype=np.uint8)
        for name, dims, ttype, toff in infos:
            n = 1
            for d in dims:
                n *= d
            nbytes = tensor_nbytes(ttype, n)
            a = self.data_start + toff
            if a + nbytes > len(self._buf):
                raise ValueError(f"{self.path}: tensor {name} runs past end of file")
            self.tensors[name] = GGUFTensor(name, ttype, dims, a, nbytes, whole[a:a + nbytes])

    def close(self):
        self.tensors.clear()
        self._buf.release()
        try:
            self._mm.close()
        except BufferError:
            pass  # numpy views still alive; the OS unmaps at process exit
        self._f.close()


# ---------------------------------------------------------------------------
# writer (used by the synthetic-model generator and the tests)
# ---------------------------------------------------------------------------
def _enc_str(s) -> bytes:
    b = s if isinstance(s, (bytes, bytearray)) else str(s).encode("utf-8")
    return struct.pack("<Q", len(b)) + bytes(b)


def _enc_value(v) -> bytes:
    """Encode a python value with an inferred GGUF type (type tag included)."""
    if isinstance(v, bool):
        return struct.pack("<I?", _BOOL, v)
    if isinstance(v, int):
        if v < 0:
            return struct.pack("<Ii", _I32, v)
        return struct.pack("<II", _U32, v) if v < 2 ** 32 else struct.pack("<IQ", _U64, v)
    if isinstance(v, float):
        return struct.pack("<If", _F32, v)
    if isinstance(v, (str, bytes)):
        return struct.pack("<I", _STR) + _enc_str(v)
    if isinstance(v, np.ndarray):
        rev = {np.dtype(t): k for k, t in _NP_FMT.items()}
        it = rev[v.dtype]
        return struct.pack("<IIQ", _ARR, it, v.size) + v.astype(v.dtype.newbyteorder("<")).tobytes()
    if isinstance(v, (list, tuple)):
        if len(v) and isinstance(v[0], (str, bytes)):
            return struct.pack("<IIQ", _ARR, _STR, len(v)) + b"".join(_enc_str(x) for x in v)
        if len(v) and isinstance(v[0], float):
            return _enc_value(np.asarray(v, dtype=np.float32))
        return _enc_value(np.asarray(v, dtype=np.int32))
    raise TypeError(f"cannot encode {type(v)} as a GGUF value")


def write_gguf(path: str, metadata: Dict[str, Any],
               tensors: Sequence[Tuple[str, int, Tuple[int, ...], Any]],
               alignment: int = GGUF_DEFAULT_ALIGNMENT) -> None:
    """Write a GGUF v3 file.

    ``tensors``: (name, ggml_type, shape in ggml order [K, N...], raw block bytes).
    The bytes may be a numpy uint8 array or a callable ``fn() -> np.ndarray`` so a
    multi-GB synthetic model never has to sit in RAM twice.
    """
    head = bytearray()
    head += struct.pack("<IIQQ", GGUF_MAGIC, 3, len(tensors), len(metadata))
    for k, v in metadata.items():
        head += _enc_str(k) + _enc_value(v)
    off = 0
    sizes = []
    for name, ttype, shape, _ in tensors:
        n = 1
        for d in shape:
            n *= d
        nbytes = tensor_nbytes(ttype, n)
        head += _enc_str(name) + struct.pack("<I", len(shape))
        head += b"".join(struct.pack("<Q", d) for d in shape)
        head += struct.pack("<IQ", ttype, off)
        sizes.append(nbytes)
        off = (off + nbytes + alignment - 1) // alignment * alignment
    pad = (-len(head)) % alignment
    with open(path, "wb") as f:
        f.write(head)
        f.write(b"\0" * pad)
        for (name, _, _, data), nbytes in zip(tensors, sizes):
            arr = data() if callable(data) else data
            arr = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
            if arr.size != nbytes:
                raise ValueError(f"tensor {name}: got {arr.size} bytes, expected {nbytes}")
            f.write(arr.data)
            f.write(b"\0" * ((-nbytes) % alignment))
