"""The serving launcher / replica router / health route (llama_cpp/server.py) on a fake engine (CPU):
defaults reproduce the reference's behaviour (compared with the unmodified /root/reference/api.py where that checkout
exists), raised limits let concurrent requests reach the continuous-batching scheduler, a 408 stops the generation,
the router spreads load and skips unhealthy replicas; real incremental streaming on the Llama object."""
import importlib.util
import os
import sys
import threading
import time

import pytest

from fake_engine import FakeEngine

REF = "/root/reference/api.py"
BODY = {"bot_profile": {"name": "Ann.f", "appearance": "a,b,c,tall,green eyes", "system_prompt": ""},
        "user_profile": {"name": "u"},
        "context": [{"turn": "user", "message": "hello world"}, {"turn": "bot", "message": "hi"},
                    {"turn": "user", "message": "what time is it now?"}]}


@pytest.fixture()
def model(tmp_path, monkeypatch):
    from llama_cpp import synth
    import llama_cpp.llama as L
    monkeypatch.setattr(L, "Engine", FakeEngine)
    FakeEngine.delay, FakeEngine.eos_after, FakeEngine.fail_with = 0.0, 12, None
    path = str(tmp_path / "m.gguf")
    synth.write_synthetic_model(path, "tiny", "Q4_K_M", seed=1)
    return path


def _llm(path, n_seq_max=1):
    from llama_cpp import Llama
    llm = Llama(model_path=path, n_gpu_layers=-1, n_ctx=1024, n_seq_max=n_seq_max, verbose=False)
    FakeEngine.eos_token = llm._vocab.eot_id
    return llm


def _many(client, n, body=BODY):
    codes = []

    def post():
        codes.append(client.post("/response", json=body).status_code)
    th = [threading.Thread(target=post) for _ in range(n)]
    [t.start() for t in th]
    [t.join(120) for t in th]
    return codes


def test_defaults_reproduce_the_reference_app(model, tmp_path, monkeypatch):
    from fastapi.testclient import TestClient
    from llama_cpp import server
    llm = _llm(model)
    with TestClient(server.create_app([llm])) as client:
        r = client.post("/response", json=BODY)
        assert r.status_code == 200 and isinstance(r.json()["response"], str)
        ours = r.json()["response"]
        assert client.post("/response", json={"bot_profile": {}}).status_code == 422
        assert client.get("/items/7").json() == {"item_id": 7}
        h = client.get("/health")
        assert h.status_code == 200 and h.json()["status"] == "ok"
        assert h.json()["limits"] == {"queue": 5, "permits": 1, "consumers": 1, "timeout_s": 25}
    llm.close()
    if not os.path.exists(REF):
        return
    # the unmodified reference app on the same fake engine answers the same request with the same text: the prompt
    # assembly, truncation and sampling settings are the reference's
    work = tmp_path / "app"
    (work / "models").mkdir(parents=True)
    os.symlink(model, work / "models" / "Lexi-Llama-3-8B-Uncensored_Q4_K_M.gguf")
    monkeypatch.chdir(work)
    monkeypatch.syspath_prepend("/root/reference")
    for m in ("ref_api_for_server_test", "data", "data.requests"):
        sys.modules.pop(m, None)
    spec = importlib.util.spec_from_file_location("ref_api_for_server_test", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    with TestClient(mod.app) as client:
        theirs = client.post("/response", json=BODY).json()["response"]
    mod.llm.close()
    assert ours == theirs


def test_reference_back_pressure_with_default_limits(model):
    from fastapi.testclient import TestClient
    from llama_cpp import server
    llm = _llm(model)
    FakeEngine.delay, FakeEngine.eos_after = 0.05, None
    with TestClient(server.create_app([llm], timeout_seconds=0.6)) as client:
        codes = _many(client, 10)
    assert 503 in codes and 408 in codes and set(codes) <= {200, 408, 503}       # 1 running + 5 queued, the rest refused
    FakeEngine.delay = 0.0
    llm.close()


def test_32_concurrent_requests_are_batched(model):
    from fastapi.testclient import TestClient
    from llama_cpp import server
    llm = _llm(model, n_seq_max=32)
    FakeEngine.delay, FakeEngine.eos_after = 0.01, 24
    with TestClient(server.create_app([llm], max_queue_size=64, permits=32, consumers=32)) as client:
        codes = _many(client, 32)
        h = client.get("/health").json()
    assert codes == [200] * 32
    assert llm._scheduler.stats["max_batch"] >= 16 and llm._engine.calls["max_batch"] >= 16
    assert h["replicas"][0]["max_running"] >= 16 and h["replicas"][0]["served"] == 32
    FakeEngine.delay = 0.0
    llm.close()


def test_timeout_cancels_the_generation(model):
    from fastapi.testclient import TestClient
    from llama_cpp import server
    llm = _llm(model)
    FakeEngine.delay, FakeEngine.eos_after = 0.02, None       # would run for ~1000 tokens = 20 s
    with TestClient(server.create_app([llm], timeout_seconds=0.3)) as client:
        t0 = time.time()
        assert client.post("/response", json=BODY).status_code == 408
        for _ in range(100):               # the scheduler notices the cancellation at its next step
            if llm._scheduler.stats["cancelled"]:
                break
            time.sleep(0.02)
        # the engine is free again right away: the next request is served (a reference server would still be busy for ~20 s)
        FakeEngine.delay, FakeEngine.eos_after = 0.0, 6
        assert client.post("/response", json=BODY).status_code == 200
        assert time.time() - t0 < 5.0
    assert llm._scheduler.stats["cancelled"] >= 1
    llm.close()


def test_router_spreads_load_and_skips_unhealthy_replicas(model):
    from fastapi.testclient import TestClient
    from llama_cpp import server
    a, b = _llm(model, n_seq_max=4), _llm(model, n_seq_max=4)
    FakeEngine.delay, FakeEngine.eos_after = 0.01, 16
    with TestClient(server.create_app([a, b], max_queue_size=8, permits=4, consumers=4)) as client:
        assert _many(client, 12) == [200] * 12
        reps = client.get("/health").json()["replicas"]
        assert reps[0]["served"] + reps[1]["served"] == 12 and min(reps[0]["served"], reps[1]["served"]) >= 3
        # replica 0 starts failing: after three errors in a row it is taken out, the service stays up on replica 1
        FakeEngine.delay = 0.0
        router = client.app.state.router
        router.replicas[0].errors_in_a_row = 3
        assert _many(client, 4) == [200] * 4
        assert client.get("/health").json()["replicas"][0]["healthy"] is False
        router.replicas[1].errors_in_a_row = 3
        assert client.post("/response", json=BODY).status_code == 503 and client.get("/health").status_code == 503
    a.close(); b.close()


def test_streaming_is_incremental_and_cancellable(model):
    llm = _llm(model)
    FakeEngine.delay, FakeEngine.eos_after = 0.01, 20
    prompt = [5, 6, 7, 8]
    whole = llm.create_completion(prompt, max_tokens=32, temperature=0.0, repeat_penalty=1.0)["choices"][0]
    t0 = time.time()
    stamps, text, finish = [], "", None
    for ch in llm.create_completion(prompt, max_tokens=32, temperature=0.0, repeat_penalty=1.0, stream=True):
        stamps.append(time.time() - t0)
        text += ch["choices"][0]["text"]
        finish = ch["choices"][0]["finish_reason"] or finish
    assert text == whole["text"] and finish == whole["finish_reason"]
    assert len(stamps) >= 5 and stamps[0] < 0.5 * stamps[-1]          # chunks arrive while the generation is running
    # a stop string split over tokens is never emitted, not even partly
    stop = whole["text"][6:9]
    out = "".join(c["choices"][0]["text"] for c in llm.create_completion(prompt, max_tokens=32, temperature=0.0, repeat_penalty=1.0,
                                                                         stream=True, stop=[stop]))
    assert out == whole["text"][:6]
    # walking away from the generator cancels the request
    FakeEngine.eos_after = None
    gen = llm.create_completion(prompt, max_tokens=900, temperature=0.0, repeat_penalty=1.0, stream=True)
    next(gen); next(gen)
    gen.close()
    time.sleep(0.2)
    assert llm._scheduler.stats["cancelled"] >= 1 and all(s.req is None for s in llm._scheduler.slots)
    # chat chunks keep the upstream shape
    FakeEngine.eos_after = 8
    chunks = list(llm.create_chat_completion(messages=[{"role": "user", "content": "hi"}], stream=True, temperature=0.0))
    assert chunks[0]["choices"][0]["delta"] == {"role": "assistant"} and chunks[-1]["choices"][0]["finish_reason"] in ("stop", "length")
    FakeEngine.delay = 0.0
    llm.close()
