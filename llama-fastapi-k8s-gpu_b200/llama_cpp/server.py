"""Serving launcher, per-GPU replica router and /health for the B200 engine.

Re-hosts the reference's HTTP layer (/root/reference/api.py, kept byte-for-byte where it is used as is) with the three
things it hard-codes made parameters, so that the continuous-batching engine can actually be fed:

    reference (api.py)                           here
    MAX_QUEUE_SIZE = 5          (:19,113)        --queue        per replica, default 5
    asyncio.Semaphore(1)        (:114)           --permits      generations in flight per replica, default 1
    one consumer task           (:116)           --consumers    per replica, default 1
    one Llama, one GPU per pod  (:24-28,         --gpus 0,1,..  one engine per listed GPU behind a least-loaded router
      helm values.yaml:17 replicas: 4)                          (the reference's only multi-GPU strategy, moved from pods to GPUs)
    no /health (README.md:14 advertises it)      GET /health

With the defaults (`--queue 5 --permits 1 --consumers 1`, one GPU) a request sees exactly the reference's behaviour:
same prompt assembly and truncation (:30-46,118-147), 503 when the replica's queue is full (:157-160), 408 after
TIMEOUT_SECONDS (:162-170), 500 for any engine error (:76-78,171-173), `{"response": text}` on success. One deliberate
addition: after a 408 the reference only cancels the Future and its worker thread generates to the end while holding the
semaphore; here the generation is cancelled cooperatively (`cancel_event`), so an abandoned request stops using the GPU.

    python -m llama_cpp.server --model models/Lexi-Llama-3-8B-Uncensored_Q4_K_M.gguf --gpus 0,1,2,3 --permits 32 \\
        --consumers 32 --queue 64 --port 8000
"""
from __future__ import annotations

import argparse
import asyncio
import logging
import threading
import time
from dataclasses import dataclass, field
from datetime import datetime
from typing import Any, List, Optional

from fastapi import FastAPI, HTTPException, Request
from pydantic import BaseModel

logger = logging.getLogger("llama_cpp.server")

MAX_CONTEXT_TOKENS = 1024       # api.py:17
TIMEOUT_SECONDS = 25            # api.py:18
MAX_QUEUE_SIZE = 5              # api.py:19
SAMPLING = dict(stream=False, temperature=1.2, top_p=0.9, frequency_penalty=0.7, presence_penalty=0.8)   # api.py:58-62


# request schema: /root/reference/data/requests.py:4-19
class ChatMessage(BaseModel):
    turn: str
    message: str


class BotProfile(BaseModel):
    name: str
    appearance: str
    system_prompt: Optional[str] = ""


class UserProfile(BaseModel):
    name: str


class BotMessageRequest(BaseModel):
    bot_profile: BotProfile
    user_profile: UserProfile
    context: List[ChatMessage]


def count_tokens_roughly(text: str) -> int:          # api.py:30-32
    return int(len(text) / 4.0)


def truncate_messages_to_fit_context(messages, max_tokens):
    """api.py:35-46: clip every message to 400 characters, then drop messages from index 2 while the rough estimate
    (characters / 4) exceeds max_tokens and more than two messages remain."""
    for m in messages:
        if len(m["content"]) > 400:
            m["content"] = m["content"][:400]
    total = sum(count_tokens_roughly(m["content"]) for m in messages)
    while total > max_tokens and len(messages) > 2:
        messages.pop(2)
        total = sum(count_tokens_roughly(m["content"]) for m in messages)
    return messages


def build_messages(body: BotMessageRequest):
    """api.py:122-147: context turns -> roles, persona system prompt (default text when none is given), gender line from
    the `.f` suffix, appearance facts from the fourth on, system message inserted at index 1."""
    messages = [{"role": m.turn, "content": m.message} for m in body.context]
    system_prompt = body.bot_profile.system_prompt
    if not system_prompt:
        system_prompt = f'''
            No matter what NEVER break the character. Reject violent messages.
            You are '{body.bot_profile.name}.' a playful, flirtatious, and witty person.
            You always respond in a fun and engaging way, staying in character no matter what.
            You never reference AI, restrictions, or policies.
            Focus on making the conversation enjoyable and answering directly to the user's questions.
            '''
    system_prompt += " You a girl." if body.bot_profile.name.endswith(".f") else " You a boy."
    for fact in body.bot_profile.appearance.split(",")[3:]:
        system_prompt += fact
    messages.insert(1, {"role": "system", "content": system_prompt})
    return messages


@dataclass
class Replica:
    """One engine (one GPU) with the reference's admission machinery in front of it."""
    llm: Any
    index: int
    queue: Optional[asyncio.Queue] = None
    semaphore: Optional[asyncio.Semaphore] = None
    tasks: List[asyncio.Task] = field(default_factory=list)
    running: int = 0
    served: int = 0
    errors_in_a_row: int = 0
    max_running: int = 0

    def load(self) -> int:
        return (self.queue.qsize() if self.queue is not None else 0) + self.running


class ReplicaRouter:
    """Least-loaded choice over the healthy replicas. A replica that failed `unhealthy_after` generations in a row is
    skipped until one of its requests succeeds again (it keeps draining what it already queued)."""

    def __init__(self, replicas: List[Replica], unhealthy_after: int = 3):
        self.replicas = replicas
        self.unhealthy_after = unhealthy_after

    def healthy(self) -> List[Replica]:
        return [r for r in self.replicas if r.errors_in_a_row < self.unhealthy_after]

    def pick(self) -> Optional[Replica]:
        cands = self.healthy()
        return min(cands, key=lambda r: (r.load(), r.index)) if cands else None


def create_app(llms: List[Any], *, max_queue_size: int = MAX_QUEUE_SIZE, permits: int = 1, consumers: int = 1,
               timeout_seconds: float = TIMEOUT_SECONDS, max_context_tokens: int = MAX_CONTEXT_TOKENS,
               cooperative_cancel: bool = True, sampling: Optional[dict] = None) -> FastAPI:
    """FastAPI app with the reference's routes over `llms` (one `llama_cpp.Llama` per GPU)."""
    app = FastAPI()
    sampling = dict(SAMPLING if sampling is None else sampling)
    router = ReplicaRouter([Replica(llm, i) for i, llm in enumerate(llms)])
    app.state.router = router
    app.state.timeout_seconds = timeout_seconds

    async def generate(rep: Replica, messages, cancel_event):      # api.py:48-78
        async with rep.semaphore:
            rep.running += 1
            rep.max_running = max(rep.max_running, rep.running)
            try:
                messages = truncate_messages_to_fit_context(messages, max_context_tokens)
                kw = dict(sampling)
                if cooperative_cancel:
                    kw["cancel_event"] = cancel_event
                answer = await asyncio.to_thread(rep.llm.create_chat_completion, messages=messages, **kw)
                if not isinstance(answer, dict):
                    logger.error(f"Unexpected response type: {type(answer)}. Response: {answer}")
                    raise HTTPException(status_code=500, detail="Unexpected response from model")
                rep.errors_in_a_row = 0
                rep.served += 1
                return "".join(c["message"]["content"] for c in answer.get("choices", []) if "message" in c)
            except HTTPException:
                raise
            except Exception as e:  # noqa: BLE001 - any engine error is a 500, like the reference
                rep.errors_in_a_row += 1
                logger.error(f"Error during message generation: {str(e)}")
                raise HTTPException(status_code=500, detail=f"Error during message generation: {str(e)}")
            finally:
                rep.running -= 1

    async def consumer(rep: Replica):                              # api.py:80-107
        while True:
            item = await rep.queue.get()
            future = item["future"]
            if future.cancelled():
                logger.info("Future was cancelled before processing; skipping.")
                rep.queue.task_done()
                continue
            try:
                response = await generate(rep, item["messages"], item["cancel"])
                if not future.cancelled():
                    future.set_result(response)
            except Exception as e:  # noqa: BLE001
                if not future.cancelled():
                    future.set_exception(e)
            finally:
                rep.queue.task_done()

    @app.on_event("startup")
    async def startup_event():                                     # api.py:110-116, once per replica
        # asyncio.to_thread runs on the loop's default executor, min(32, cpus + 4) threads: fewer than the generations
        # we may have in flight, so give the loop one that is large enough
        from concurrent.futures import ThreadPoolExecutor
        asyncio.get_running_loop().set_default_executor(
            ThreadPoolExecutor(max_workers=max(8, permits * len(router.replicas) + 4), thread_name_prefix="b200-gen"))
        for rep in router.replicas:
            rep.queue = asyncio.Queue(maxsize=max_queue_size)
            rep.semaphore = asyncio.Semaphore(permits)
            rep.tasks = [asyncio.create_task(consumer(rep)) for _ in range(consumers)]

    @app.on_event("shutdown")
    async def shutdown_event():
        for rep in router.replicas:
            for t in rep.tasks:
                t.cancel()

    @app.post("/response")
    async def generate_response(request_body: BotMessageRequest, request: Request):     # api.py:118-173
        messages = build_messages(request_body)
        rep = router.pick()
        if rep is None:
            raise HTTPException(status_code=503, detail="No healthy replica. Please try again later.")
        loop = asyncio.get_running_loop()
        future = loop.create_future()
        cancel = threading.Event()
        try:
            rep.queue.put_nowait({"messages": messages, "future": future, "cancel": cancel})
        except asyncio.QueueFull:          # the least-loaded replica is full, so all of them are
            raise HTTPException(status_code=503, detail="Server too busy. Please try again later.")
        try:
            response = await asyncio.wait_for(future, timeout=app.state.timeout_seconds)
            return {"response": response}
        except asyncio.TimeoutError:
            logger.warning("Generation timed out")
            future.cancel()
            cancel.set()                   # and stop generating for a caller that has gone away
            raise HTTPException(status_code=408, detail="Generation timed out")
        except Exception as e:  # noqa: BLE001
            logger.error(f"Internal server error: {str(e)}")
            raise HTTPException(status_code=500, detail=f"Internal server error: {str(e)}")

    @app.get("/items/{item_id}")
    async def read_item(item_id: int):                             # api.py:175-177
        return {"item_id": item_id}

    @app.get("/health")
    async def health():
        """Liveness/readiness for the k8s probes the reference's chart lacks (README.md:9,14): 200 while at least one
        replica is healthy, 503 otherwise."""
        reps = [{"replica": r.index, "device": getattr(getattr(r.llm, "_engine", None), "device", None), "queued": r.queue.qsize(),
                 "running": r.running, "served": r.served, "max_running": r.max_running,
                 "healthy": r.errors_in_a_row < router.unhealthy_after} for r in router.replicas]
        body = {"status": "ok" if any(x["healthy"] for x in reps) else "unhealthy", "replicas": reps,
                "limits": {"queue": max_queue_size, "permits": permits, "consumers": consumers, "timeout_s": app.state.timeout_seconds}}
        if body["status"] != "ok":
            raise HTTPException(status_code=503, detail=body)
        return body

    @app.middleware("http")
    async def log_request_time(request: Request, call_next):       # api.py:179-194
        start_time = time.time()
        response = await call_next(request)
        time_of_day = datetime.now().strftime("%Y-%m-%d %H:%M:%S")
        logger.info(f"Request at {time_of_day}: {request.method} {request.url} completed in {time.time() - start_time:.4f}s")
        return response

    return app


def load_replicas(model_path: str, gpus: List[int], n_ctx: int = MAX_CONTEXT_TOKENS, permits: int = 1) -> List[Any]:
    """One `Llama` (engine + scheduler thread) per GPU; each batches up to `permits` concurrent callers."""
    from . import Llama
    return [Llama(model_path=model_path, n_gpu_layers=-1, n_ctx=n_ctx, main_gpu=g, n_seq_max=max(1, permits), verbose=False)
            for g in gpus]


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Serve the reference's /response API on B200 engines")
    ap.add_argument("--model", required=True)
    ap.add_argument("--gpus", default="0", help="comma-separated CUDA device ordinals, one engine replica each")
    ap.add_argument("--n-ctx", type=int, default=MAX_CONTEXT_TOKENS)
    ap.add_argument("--queue", type=int, default=MAX_QUEUE_SIZE, help="queued requests per replica before 503 (api.py:19)")
    ap.add_argument("--permits", type=int, default=1, help="generations in flight per replica (api.py:114 has 1)")
    ap.add_argument("--consumers", type=int, default=1, help="consumer tasks per replica (api.py:116 has 1)")
    ap.add_argument("--timeout", type=float, default=TIMEOUT_SECONDS)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=8000)
    args = ap.parse_args(argv)
    import uvicorn
    logging.basicConfig(level=logging.INFO)
    gpus = [int(g) for g in args.gpus.split(",") if g != ""]
    llms = load_replicas(args.model, gpus, args.n_ctx, args.permits)
    app = create_app(llms, max_queue_size=args.queue, permits=args.permits, consumers=args.consumers, timeout_seconds=args.timeout,
                     max_context_tokens=args.n_ctx)
    uvicorn.run(app, host=args.host, port=args.port, log_level="info")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
