import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def native_lib():
    """libb200gguf.so, built in-tree if it is stale (nvcc cross-compiles without a GPU)."""
    sys.path.insert(0, PKG)
    import build as b2build
    b2build.build()
    from llama_cpp import _cabi
    return _cabi.load_library()


@pytest.fixture(scope="session")
def tmp_models(tmp_path_factory):
    """Directory for synthetic GGUF files shared by the session."""
    return str(tmp_path_factory.mktemp("models"))
