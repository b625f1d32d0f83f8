import os

import numpy as np
import pytest


def test_writer_reader_round_trip(tmp_path):
    from llama_cpp.gguf_file import GGUFFile, write_gguf
    rng = np.random.default_rng(0)
    q = rng.integers(0, 256, size=2 * 144, dtype=np.uint8)
    f = rng.standard_normal(8).astype(np.float32)
    meta = {"general.architecture": "llama", "a.int": 7, "a.neg": -3, "a.float": 1.5, "a.bool": True,
            "a.str": "héllo", "a.arr": np.arange(5, dtype=np.int32), "a.strs": [b"x", b"yz"], "a.big": 2 ** 40}
    p = str(tmp_path / "t.gguf")
    write_gguf(p, meta, [("w", 12, (256, 2), q), ("n", 0, (8,), f.view(np.uint8))])
    g = GGUFFile(p)
    assert g.metadata["a.int"] == 7 and g.metadata["a.neg"] == -3 and g.metadata["a.bool"] is True
    assert g.metadata["a.str"] == "héllo" and g.metadata["a.big"] == 2 ** 40
    assert abs(g.metadata["a.float"] - 1.5) < 1e-7
    assert list(g.metadata["a.arr"]) == [0, 1, 2, 3, 4] and g.metadata["a.strs"] == [b"x", b"yz"]
    assert g.tensors["w"].shape == (256, 2) and g.tensors["w"].type_name == "Q4_K"
    assert np.array_equal(g.tensors["w"].data, q) and np.array_equal(g.tensors["n"].data.view(np.float32), f)
    assert g.tensors["w"].offset % 32 == 0 and g.tensors["n"].offset % 32 == 0


def test_reader_agrees_with_gguf_py(tmp_path):
    gguf = pytest.importorskip("gguf")
    from llama_cpp import synth
    from llama_cpp.gguf_file import GGUFFile
    p = str(tmp_path / "m.gguf")
    synth.write_synthetic_model(p, "tiny", "Q4_K_M", seed=3)
    mine = GGUFFile(p)
    theirs = gguf.GGUFReader(p)
    assert len(theirs.tensors) == len(mine.tensors)
    for t in theirs.tensors:
        m = mine.tensors[t.name]
        assert t.tensor_type.name == m.type_name and t.data_offset == m.offset
        assert tuple(int(x) for x in t.shape) == m.shape
        assert np.array_equal(np.asarray(t.data).view(np.uint8).reshape(-1), m.data)
    assert int(theirs.fields["llama.block_count"].parts[-1][0]) == mine.metadata["llama.block_count"]


def test_bad_files_are_rejected(tmp_path):
    from llama_cpp.gguf_file import GGUFFile, write_gguf
    p = tmp_path / "bad.gguf"
    p.write_bytes(b"NOPE" + b"\0" * 64)
    with pytest.raises(ValueError, match="not a GGUF"):
        GGUFFile(str(p))
    with pytest.raises(ValueError):
        write_gguf(str(tmp_path / "x.gguf"), {}, [("w", 12, (100,), np.zeros(10, np.uint8))])   # ragged block count
    good = str(tmp_path / "g.gguf")
    write_gguf(good, {}, [("w", 0, (64,), np.zeros(256, np.uint8))])
    data = open(good, "rb").read()
    trunc = tmp_path / "trunc.gguf"
    trunc.write_bytes(data[:-100])
    with pytest.raises(ValueError, match="past end"):
        GGUFFile(str(trunc))


def test_q4_k_m_type_mix_matches_survey():
    from llama_cpp import synth
    a = synth.ARCHS["llama-3-8b"]
    types = synth.tensor_types(a, "Q4_K_M")
    shapes = synth.tensor_shapes(a)
    from llama_cpp.gguf_file import GGML_TYPE_ID, tensor_nbytes
    total = sum(tensor_nbytes(GGML_TYPE_ID[types[n]], int(np.prod(s))) for n, s in shapes.items())
    assert abs(total / 1e9 - 4.913) < 0.01            # 4.913 GB
    more = [i for i in range(32) if types[f"blk.{i}.ffn_down.weight"] == "Q6_K"]
    assert more == [0, 1, 2, 3, 6, 9, 12, 15, 18, 21, 24, 27, 28, 29, 30, 31]
    w_read = total - tensor_nbytes(GGML_TYPE_ID[types["token_embd.weight"]], a.n_embd * a.n_vocab)
    assert abs(w_read / 1e9 - 4.617) < 0.01           # W_read 4.6174 GB
