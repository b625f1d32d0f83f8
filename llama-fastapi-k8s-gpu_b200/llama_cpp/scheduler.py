"""Continuous-batching scheduler: many concurrent `create_chat_completion` callers, one engine.

The reference admits one generation at a time (asyncio.Queue(5) -> single consumer ->
Semaphore(1), /root/reference/api.py:80-116) and upstream's `Llama` is not re-entrant. This
engine is: callers on any thread (api.py:55 uses asyncio.to_thread) enqueue a Request and block;
one scheduler thread owns the engine, admits requests into free sequence slots (prefill, with
longest-common-prefix reuse of the slot's previous KV -- upstream's H5 behaviour), then advances
all live sequences one token per iteration. With a single live sequence and nobody waiting it
decodes in run-ahead chunks (the sampled token never leaves the GPU between steps) and rewinds
if a stop condition fell inside the chunk.

Stop handling restates llama-cpp-python 0.2.77 `_create_completion`: an
end-of-generation token ends with finish_reason "stop" and is not emitted; a stop string cuts the
text at its first occurrence; reaching max_tokens ends with "length".
"""
from __future__ import annotations

import threading
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Callable, Deque, List, Optional

import numpy as np

from .engine import Engine, SamplingSettings
from .tokenizer import Vocab


@dataclass
class Request:
    prompt: List[int]
    max_tokens: int
    sampling: SamplingSettings
    stop: List[bytes] = field(default_factory=list)
    uniforms: Optional[np.ndarray] = None
    on_token: Optional[Callable[[int], None]] = None     # streaming hook (scheduler thread)
    cancel_event: Optional[threading.Event] = None       # set by the caller to abandon the request (e.g. after a 408)
    # results
    completion_tokens: List[int] = field(default_factory=list)
    raw: bytearray = field(default_factory=bytearray)     # bytes of the completion so far (stop-string search window)
    text: bytes = b""
    finish_reason: Optional[str] = None
    error: Optional[BaseException] = None
    t_submit: float = 0.0
    t_first: float = 0.0
    t_done: float = 0.0
    n_reused: int = 0
    done: threading.Event = field(default_factory=threading.Event)

    def cancelled(self) -> bool:
        return self.cancel_event is not None and self.cancel_event.is_set()

    def wait(self):
        self.done.wait()
        if self.error is not None:
            raise self.error
        return self


class _Slot:
    def __init__(self, idx: int):
        self.idx = idx
        self.req: Optional[Request] = None
        self.cached: List[int] = []       # tokens whose K/V are in this slot's pages
        self.pending: int = -1            # sampled, not yet forwarded


def check_token(req: Request, vocab: Vocab, tok: int) -> bool:
    """Account for one sampled token; returns True when the request is finished. The bytes generated so far are kept in
    `req.raw` and only their tail is searched for stop strings: re-detokenising the whole completion for every token made
    a batched step of 32 long sequences cost more on the host than on the GPU."""
    if vocab.is_eog(tok):
        req.text = vocab.detokenize(req.completion_tokens)
        req.finish_reason = "stop"
        return True
    req.completion_tokens.append(tok)
    if req.on_token is not None:
        req.on_token(tok)
    if req.stop:
        piece = vocab.token_to_piece(tok, False) if vocab.model != "none" else b"<%d>" % tok
        req.raw += piece
        reach = len(piece) + max(len(s) for s in req.stop) - 1
        tail = bytes(req.raw[-reach:]) if reach < len(req.raw) else bytes(req.raw)
        if any(s in tail for s in req.stop):
            all_text = vocab.detokenize(req.completion_tokens)
            hits = [s for s in req.stop if s in all_text]
            if hits:
                req.text = all_text[:all_text.index(hits[0])]
                req.finish_reason = "stop"
                return True
    if len(req.completion_tokens) >= req.max_tokens:
        req.text = vocab.detokenize(req.completion_tokens)
        req.finish_reason = "length"
        return True
    return False


class Scheduler:
    RUN_AHEAD = 8
    ADMIT_PER_STEP = 2     # while sequences are decoding, at most this many new prompts are processed between two decode steps

    def __init__(self, engine: Engine, vocab: Vocab):
        self.engine = engine
        self.vocab = vocab
        self.slots = [_Slot(i) for i in range(engine.max_seqs)]
        self.queue: Deque[Request] = deque()
        self.cv = threading.Condition()
        self.thread: Optional[threading.Thread] = None
        self.stopping = False
        self.stats = dict(steps=0, tokens=0, max_batch=0, prefill_tokens=0, reused_tokens=0, cancelled=0,
                          t_admit=0.0, t_decode=0.0, t_host=0.0)     # seconds in prompt passes / engine decode calls / host bookkeeping

    # ------------------------------------------------------------------ public
    def submit(self, req: Request) -> Request:
        req.t_submit = time.perf_counter()
        self.stats["last_submit"] = req.t_submit
        if not self.stats.get("first_submit"):
            self.stats["first_submit"] = req.t_submit
        with self.cv:
            if self.stopping:
                raise RuntimeError("engine is shut down")
            self.queue.append(req)
            if self.thread is None:
                self.thread = threading.Thread(target=self._run, name="b200-scheduler", daemon=True)
                self.thread.start()
            self.cv.notify_all()
        return req

    def shutdown(self):
        with self.cv:
            self.stopping = True
            self.cv.notify_all()
        if self.thread is not None:
            self.thread.join(timeout=30)

    def cancel(self, req: Request):
        """Abandon a request: still queued -> it never starts; running -> its slot is freed at the next step (the caller's
        `wait()` returns with finish_reason "cancelled" and the text so far). The reference cannot do this: after a 408 its
        worker thread generates to the end while holding the semaphore (/root/reference/api.py:162-170)."""
        if req.cancel_event is None:
            req.cancel_event = threading.Event()
        req.cancel_event.set()
        with self.cv:
            self.cv.notify_all()

    def drop_cache(self):
        """Forget cached prefixes (used by Llama.reset / direct eval)."""
        for s in self.slots:
            if s.req is None:
                s.cached = []

    # ------------------------------------------------------------------ loop
    def _run(self):
        while True:
            with self.cv:
                while not self.queue and not any(s.req for s in self.slots) and not self.stopping:
                    self.cv.wait()
                if self.stopping and not self.queue and not any(s.req for s in self.slots):
                    return
                admitted = []
                n_free = sum(1 for s in self.slots if s.req is None)
                n_live = len(self.slots) - n_free
                # prompts are interleaved with decode steps: a burst of arrivals does not stall the running sequences
                # behind all of its prompt passes
                limit = n_free if n_live == 0 else min(n_free, self.ADMIT_PER_STEP)
                while self.queue and len(admitted) < limit:
                    req = self.queue.popleft()
                    if req.cancelled():
                        self._cancelled(req)
                        continue
                    admitted.append(req)
                waiting = len(self.queue)
            with self.engine.lock:
                t0 = time.perf_counter()
                for s in self.slots:
                    if s.req is not None and s.req.cancelled():
                        self._cancelled(s.req)
                        s.req = None
                for req in admitted:
                    self._admit(req)
                t1 = time.perf_counter()
                d0 = self.stats["t_decode"]
                self._step(waiting)
                self.stats["t_admit"] += t1 - t0
                self.stats["t_host"] += (time.perf_counter() - t1) - (self.stats["t_decode"] - d0)

    def _fail(self, slot: _Slot, exc: BaseException):
        req = slot.req
        slot.req = None
        slot.cached = []
        if req is not None:
            req.error = exc
            req.t_done = time.perf_counter()
            req.done.set()

    def _cancelled(self, req: Request):
        req.text = self.vocab.detokenize(req.completion_tokens)
        req.finish_reason = "cancelled"
        req.t_done = time.perf_counter()
        self.stats["cancelled"] += 1
        req.done.set()

    def _finish(self, slot: _Slot):
        req = slot.req
        slot.req = None
        req.t_done = time.perf_counter()
        self.stats["last_done"] = req.t_done
        req.done.set()

    def _admit(self, req: Request):
        free = [s for s in self.slots if s.req is None]

        def lcp(a, b):
            n = 0
            for x, y in zip(a, b):
                if x != y:
                    break
                n += 1
            return n

        slot = max(free, key=lambda s: lcp(s.cached, req.prompt))
        slot.req = req
        eng = self.engine
        try:
            n_keep = min(lcp(slot.cached, req.prompt), len(req.prompt) - 1)
            req.n_reused = n_keep
            eng.ensure_pages(slot.idx, min(eng.n_ctx, len(req.prompt) + req.max_tokens + 1))
            eng.seq_begin(slot.idx, n_keep, req.prompt[:n_keep], req.sampling, req.uniforms)
            eng.prefill(slot.idx, req.prompt[n_keep:])
            slot.cached = list(req.prompt)
            self.stats["prefill_tokens"] += len(req.prompt) - n_keep
            self.stats["reused_tokens"] += n_keep
            tok = eng.sample(slot.idx)
            req.t_first = time.perf_counter()
            slot.pending = tok
            if check_token(req, self.vocab, tok):
                self._finish(slot)
        except BaseException as exc:  # noqa: BLE001 - forwarded to the caller
            self._fail(slot, exc)

    def _step(self, waiting: int):
        live = [s for s in self.slots if s.req is not None]
        if not live:
            return
        eng = self.engine
        self.stats["max_batch"] = max(self.stats["max_batch"], len(live))
        try:
            if len(live) == 1 and waiting == 0:
                s = live[0]
                req = s.req
                room = eng.n_ctx - len(s.cached)
                n = max(1, min(self.RUN_AHEAD, req.max_tokens - len(req.completion_tokens), room))
                base = list(s.cached)
                td = time.perf_counter()
                toks = eng.decode_n(s.idx, n)
                self.stats["t_decode"] += time.perf_counter() - td
                self.stats["steps"] += n
                forwarded = [s.pending] + toks[:-1]          # tokens whose K/V were just written
                for i, t in enumerate(toks):
                    base.append(forwarded[i])
                    self.stats["tokens"] += 1
                    if check_token(req, self.vocab, t):
                        s.cached = base
                        if i + 1 < n:                         # steps past the stop were speculative
                            eng.rewind(s.idx, len(base), t if t >= 0 else 0, base)
                        self._finish(s)
                        return
                s.cached = base
                s.pending = toks[-1]
            else:
                td = time.perf_counter()
                toks = eng.decode([s.idx for s in live])
                self.stats["t_decode"] += time.perf_counter() - td
                self.stats["steps"] += 1
                for s, t in zip(live, toks):
                    s.cached.append(s.pending)
                    s.pending = t
                    self.stats["tokens"] += 1
                    if check_token(s.req, self.vocab, t):
                        self._finish(s)
        except BaseException as exc:  # noqa: BLE001
            for s in live:
                if s.req is not None:
                    self._fail(s, exc)
