"""The reference's api.py, imported UNCHANGED from /root/reference, served by this repo's
`llama_cpp` package (fake engine on CPU): status codes and payloads of SURVEY.md 8(b).
Skipped where the reference checkout does not exist (the GPU box)."""
import importlib.util
import os
import sys
import threading
import time

import pytest

from fake_engine import FakeEngine

REF = "/root/reference/api.py"
pytestmark = pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not on this machine")

BODY = {"bot_profile": {"name": "Ann.f", "appearance": "a,b,c,tall,green eyes", "system_prompt": ""},
        "user_profile": {"name": "u"},
        "context": [{"turn": "user", "message": "hello world"}, {"turn": "bot", "message": "hi"},
                    {"turn": "user", "message": "what time is it now?"}]}


@pytest.fixture()
def ref_app(tmp_path, monkeypatch):
    from fastapi.testclient import TestClient
    from llama_cpp import synth
    import llama_cpp.llama as L
    monkeypatch.setattr(L, "Engine", FakeEngine)
    FakeEngine.delay, FakeEngine.eos_after, FakeEngine.fail_with = 0.0, 12, None
    work = tmp_path / "app"
    (work / "models").mkdir(parents=True)
    synth.write_synthetic_model(str(work / "models" / "Lexi-Llama-3-8B-Uncensored_Q4_K_M.gguf"), "tiny", "Q4_K_M", seed=1)
    monkeypatch.chdir(work)
    monkeypatch.syspath_prepend("/root/reference")
    for m in ("ref_api_under_test", "data", "data.requests"):
        sys.modules.pop(m, None)
    spec = importlib.util.spec_from_file_location("ref_api_under_test", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)                       # runs `llm = Llama(model_path=..., n_gpu_layers=-1, n_ctx=1024)`
    FakeEngine.eos_token = mod.llm._vocab.eot_id
    with TestClient(mod.app) as client:
        yield mod, client
    mod.llm.close()


def test_reference_source_is_unmodified():
    src = open(REF).read()
    assert "from llama_cpp import Llama" in src and "n_gpu_layers=-1" in src and "create_chat_completion" in src


def test_response_200_and_422(ref_app):
    mod, client = ref_app
    r = client.post("/response", json=BODY)
    assert r.status_code == 200 and isinstance(r.json()["response"], str)
    assert client.post("/response", json={"bot_profile": {}}).status_code == 422
    assert client.get("/items/7").json() == {"item_id": 7}
    assert client.get("/health").status_code == 404          # advertised in the README, not implemented
    assert mod.llm.n_ctx() == 1024 and mod.llm.chat_format == "llama-3"


def test_engine_exception_becomes_500(ref_app):
    mod, client = ref_app
    FakeEngine.fail_with = RuntimeError("b200 engine call failed (code -2): device lost")
    r = client.post("/response", json=BODY)
    assert r.status_code == 500 and "device lost" in r.json()["detail"]
    FakeEngine.fail_with = None
    assert client.post("/response", json=BODY).status_code == 200


def test_too_long_prompt_becomes_500_with_upstream_message(ref_app):
    mod, client = ref_app
    # 30 short messages survive the character clip but exceed 1024 tokens once templated
    body = dict(BODY, context=[{"turn": "user", "message": "zq " * 130} for _ in range(11)])
    r = client.post("/response", json=body)
    assert r.status_code in (200, 500)
    if r.status_code == 500:
        assert "exceed context window of 1024" in r.json()["detail"]


def test_queue_full_503_and_timeout_408(ref_app, monkeypatch):
    mod, client = ref_app
    FakeEngine.delay, FakeEngine.eos_after = 0.05, None       # slow engine: each request takes seconds
    monkeypatch.setattr(mod, "TIMEOUT_SECONDS", 0.6)
    codes = []

    def post():
        codes.append(client.post("/response", json=BODY).status_code)

    th = [threading.Thread(target=post) for _ in range(10)]
    [t.start() for t in th]
    [t.join(120) for t in th]
    assert 503 in codes                                       # more than 1 running + 5 queued
    assert 408 in codes                                       # waited longer than TIMEOUT_SECONDS
    assert set(codes) <= {200, 408, 503}
    FakeEngine.delay = 0.0
    time.sleep(0.2)
