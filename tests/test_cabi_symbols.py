"""The C-ABI library loads without a GPU and exports every symbol include/b200_gguf.h declares
(no compute calls here)."""
import os
import re

from conftest import ROOT


def _declared():
    names = []
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            text = open(os.path.join(ROOT, "include", fn)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names += re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol(native_lib):
    decl = _declared()
    assert len(decl) >= 25
    for name in decl:
        assert hasattr(native_lib, name), f"{name} declared in include/ but not exported"


def test_python_binding_covers_every_declared_symbol(native_lib):
    from llama_cpp import _cabi
    assert sorted(_cabi.SIGNATURES) == _declared()


def test_version_and_error_string(native_lib):
    assert native_lib.b2_version() >= 100
    assert native_lib.b2_engine_destroy(None) == 0
    assert native_lib.b2_engine_finalize(None) != 0
    assert b"null" in native_lib.b2_last_error()


def test_missing_library_fails_loudly(tmp_path):
    import pytest
    from llama_cpp import _cabi
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _cabi.load_library(str(tmp_path / "nope.so"))
