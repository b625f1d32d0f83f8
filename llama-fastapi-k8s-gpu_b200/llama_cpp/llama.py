"""`llama_cpp.Llama` -- drop-in for the llama-cpp-python object the reference drives.

The reference uses exactly two things (/root/reference/api.py):
    llm = Llama(model_path=file_path, n_gpu_layers=-1, n_ctx=1024)                 # :24-28
    llm.create_chat_completion(messages=..., stream=False, temperature=1.2, top_p=0.9,
                               frequency_penalty=0.7, presence_penalty=0.8) -> dict  # :55-63
and reads `answer["choices"][i]["message"]["content"]` (:65-74). Same keywords, defaults, return
shape and error behaviour as llama-cpp-python 0.2.77 (un-vendored upstream),
on top of the B200 engine: no CPU path (`n_gpu_layers=0` raises), concurrent callers are batched.
"""
from __future__ import annotations

import os
import queue
import threading
import time
import uuid
from typing import Any, Dict, Iterator, List, Optional, Sequence, Union

import numpy as np

from . import chat_format as cf
from .engine import Engine, SamplingSettings
from .gguf_file import GGUFFile
from .sampling import UniformStream
from .scheduler import Request, Scheduler
from .tokenizer import Vocab


class Llama:
    def __init__(self, model_path: str, *, n_gpu_layers: int = 0, n_ctx: int = 512, seed: int = 0xFFFFFFFF,
                 n_batch: int = 512, last_n_tokens_size: int = 64, chat_format: Optional[str] = None,
                 verbose: bool = True, main_gpu: int = 0, n_seq_max: Optional[int] = None, tp=None, **kwargs: Any):
        """Load a GGUF model onto one B200.

        model_path / n_gpu_layers / n_ctx as in upstream. n_gpu_layers=-1 (or any non-zero value)
        means "all layers on the GPU"; 0 asks for upstream's CPU path, which this engine does not
        have. n_ctx=0 takes the model's training context. Extra upstream keywords (n_threads,
        use_mmap, rope_*, flash_attn, ...) are accepted and ignored. `n_seq_max` (or env
        B200_MAX_SEQS) is how many callers can be batched concurrently (default 1, like the reference).
        `tp` = (rank, world, nccl_id) makes this object one rank of a tensor-parallel group (70B-class models; one process
        per GPU, every rank makes the same calls in the same order)."""
        if not os.path.exists(model_path):
            raise ValueError(f"Model path does not exist: {model_path}")
        if n_gpu_layers == 0:
            raise ValueError("n_gpu_layers=0 requests the CPU path; the B200 engine has no CPU fallback "
                             "(pass n_gpu_layers=-1 as /root/reference/api.py:26 does)")
        self.model_path = model_path
        self.verbose = verbose
        self._gguf = GGUFFile(model_path)
        self.metadata: Dict[str, Any] = {k: v for k, v in self._gguf.metadata.items()
                                         if isinstance(v, (str, int, float, bool))}
        dims_ctx = int(self._gguf.metadata.get("llama.context_length", 2048))
        self._n_ctx = int(n_ctx) if n_ctx else dims_ctx
        self._seed = seed
        self.last_n_tokens_size = min(int(last_n_tokens_size), 64)
        max_seqs = int(n_seq_max or os.environ.get("B200_MAX_SEQS", "1"))
        device = int(os.environ.get("B200_DEVICE", main_gpu))
        self._engine = Engine(self._gguf, n_ctx=self._n_ctx, max_seqs=max_seqs, device=device,
                              kv_block_tokens=int(os.environ.get("B200_KV_BLOCK", "64")), tp=tp)
        self._vocab = Vocab(self._gguf.metadata, self._engine.dims.n_vocab)
        self._scheduler = Scheduler(self._engine, self._vocab)
        self._stream = UniformStream(seed)
        self._stream_lock = threading.Lock()
        # chat format selection (upstream Llama.__init__)
        self.chat_format = chat_format
        self._chat_formatter = None
        md = self._gguf.metadata
        if self.chat_format is None and "tokenizer.chat_template" in md:
            guess = cf.guess_chat_format_from_gguf_metadata(md)
            if guess is not None:
                self.chat_format = guess
            else:
                v = self._vocab
                eos = v.tokens[v.eos_id].decode("utf-8", "replace") if 0 <= v.eos_id < len(v.tokens) else ""
                bos = v.tokens[v.bos_id].decode("utf-8", "replace") if 0 <= v.bos_id < len(v.tokens) else ""
                self._chat_formatter = cf.Jinja2ChatFormatter(md["tokenizer.chat_template"], eos, bos)
        if self.chat_format is None and self._chat_formatter is None:
            self.chat_format = "llama-2"
        if self.chat_format is not None and self.chat_format not in cf.FORMATS:
            raise ValueError(f"Invalid chat handler: {self.chat_format}")
        self.last_timings: Dict[str, float] = {}
        # direct eval()/sample() state (slot 0, outside the scheduler)
        self._input_ids: List[int] = []
        self._direct_open = False

    # ------------------------------------------------------------------ small accessors
    def n_ctx(self) -> int:
        return self._n_ctx

    def n_vocab(self) -> int:
        return self._engine.dims.n_vocab

    def n_embd(self) -> int:
        return self._engine.dims.n_embd

    def token_bos(self) -> int:
        return self._vocab.bos_id

    def token_eos(self) -> int:
        return self._vocab.eos_id

    def tokenize(self, text: bytes, add_bos: bool = True, special: bool = False) -> List[int]:
        return self._vocab.tokenize(text, add_bos=add_bos, special=special)

    def detokenize(self, tokens: Sequence[int], prev_tokens: Optional[Sequence[int]] = None) -> bytes:
        return self._vocab.detokenize(list(tokens))

    def close(self):
        if getattr(self, "_scheduler", None) is not None:
            self._scheduler.shutdown()
            self._scheduler = None
        if getattr(self, "_engine", None) is not None:
            self._engine.close()
            self._engine = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ low-level API (tests / parity)
    def reset(self):
        self._input_ids = []
        self._direct_open = False
        self._scheduler.drop_cache()

    def eval(self, tokens: Sequence[int]):
        """Append tokens to the (single) direct sequence; logits of the last one via `.scores`."""
        toks = [int(t) for t in tokens]
        if not toks:
            return
        with self._engine.lock:
            eng = self._engine
            if len(self._input_ids) + len(toks) > self._n_ctx:
                raise ValueError(f"Requested tokens ({len(self._input_ids) + len(toks)}) exceed context window of {self._n_ctx}")
            self._scheduler.drop_cache()
            eng.ensure_pages(0, len(self._input_ids) + len(toks))
            if not self._direct_open:
                eng.seq_begin(0, 0, [], SamplingSettings(temperature=0.0, repeat_penalty=1.0))
                self._direct_open = True
            if len(toks) > 1 or not self._input_ids:
                eng.prefill(0, toks)
            else:
                eng.set_token(0, toks[0])
                eng.forward_pending(0)
            self._input_ids.extend(toks)

    @property
    def scores(self) -> np.ndarray:
        """Logits of the last evaluated token, shape [1, n_vocab] (upstream keeps all rows)."""
        with self._engine.lock:
            return self._engine.logits(0)[None, :]

    # ------------------------------------------------------------------ completion
    def _sampling(self, temperature, top_p, top_k, min_p, repeat_penalty, frequency_penalty, presence_penalty,
                  typical_p, tfs_z, mirostat_mode, logit_bias, grammar, logits_processor) -> SamplingSettings:
        if mirostat_mode:
            raise NotImplementedError("mirostat sampling is not implemented on the device sampler")
        if logit_bias or grammar is not None or logits_processor is not None:
            raise NotImplementedError("logit_bias / grammar / logits_processor are not implemented")
        if typical_p != 1.0 or tfs_z != 1.0:
            raise NotImplementedError("typical_p / tfs_z other than 1.0 are not implemented")
        if temperature < 0:
            raise NotImplementedError("temperature < 0 (sample from the full softmax) is not implemented")
        if temperature > 0 and not (1 <= top_k <= 1024):
            raise NotImplementedError("the device sampler supports top_k in 1..1024 (upstream default 40)")
        return SamplingSettings(temperature=float(temperature), top_p=float(top_p), min_p=float(min_p),
                                top_k=int(top_k), repeat_penalty=float(repeat_penalty),
                                frequency_penalty=float(frequency_penalty), presence_penalty=float(presence_penalty),
                                penalty_last_n=self.last_n_tokens_size)

    def _make_request(self, prompt_tokens: List[int], max_tokens: Optional[int], sampling: SamplingSettings,
                      stop: List[bytes], seed: Optional[int], on_token=None, cancel_event=None) -> Request:
        if len(prompt_tokens) == 0:
            prompt_tokens = [self._vocab.bos_id if self._vocab.bos_id >= 0 else 0]
        if len(prompt_tokens) >= self._n_ctx:
            raise ValueError(f"Requested tokens ({len(prompt_tokens)}) exceed context window of {self._n_ctx}")
        if max_tokens is None or max_tokens <= 0:
            max_tokens = self._n_ctx - len(prompt_tokens)
        if max_tokens + len(prompt_tokens) >= self._n_ctx:
            max_tokens = self._n_ctx - len(prompt_tokens)
        uniforms = None
        if sampling.temperature > 0:
            if seed is not None:
                uniforms = UniformStream(seed).take(min(max_tokens + 1, 8192))
            else:
                with self._stream_lock:       # concurrent callers share the object's generator
                    uniforms = self._stream.take(min(max_tokens + 1, 8192))
        self._direct_open = False
        self._input_ids = []
        return Request(prompt=list(prompt_tokens), max_tokens=int(max_tokens), sampling=sampling, stop=stop,
                       uniforms=uniforms, on_token=on_token, cancel_event=cancel_event)

    def _run(self, prompt_tokens: List[int], max_tokens: Optional[int], sampling: SamplingSettings,
             stop: List[bytes], seed: Optional[int], on_token=None, cancel_event=None) -> Request:
        req = self._make_request(prompt_tokens, max_tokens, sampling, stop, seed, on_token, cancel_event)
        self._scheduler.submit(req).wait()
        n = len(req.completion_tokens)
        self.last_timings = dict(
            prompt_tokens=len(req.prompt), completion_tokens=n, reused_tokens=req.n_reused,
            ttft_ms=(req.t_first - req.t_submit) * 1e3, decode_ms=(req.t_done - req.t_first) * 1e3,
            total_ms=(req.t_done - req.t_submit) * 1e3)
        return req

    def create_completion(self, prompt: Union[str, List[int]], suffix: Optional[str] = None,
                          max_tokens: Optional[int] = 16, temperature: float = 0.8, top_p: float = 0.95,
                          min_p: float = 0.05, typical_p: float = 1.0, logprobs: Optional[int] = None, echo: bool = False,
                          stop: Optional[Union[str, List[str]]] = None, frequency_penalty: float = 0.0,
                          presence_penalty: float = 0.0, repeat_penalty: float = 1.1, top_k: int = 40,
                          stream: bool = False, seed: Optional[int] = None, tfs_z: float = 1.0, mirostat_mode: int = 0,
                          mirostat_tau: float = 5.0, mirostat_eta: float = 0.1, model: Optional[str] = None,
                          stopping_criteria=None, logits_processor=None, grammar=None,
                          logit_bias: Optional[Dict[str, float]] = None, _add_bos: bool = True,
                          _special: bool = False, cancel_event: Optional[threading.Event] = None, **kwargs: Any):
        """`cancel_event` (not in upstream): set it from another thread to abandon the generation; the call then returns
        what was produced so far with finish_reason "cancelled" and the sequence slot is free again."""
        if logprobs is not None:
            raise NotImplementedError("logprobs are not implemented")
        if stopping_criteria is not None:
            raise NotImplementedError("stopping_criteria is not implemented")
        if isinstance(prompt, str):
            tokens = self.tokenize(prompt.encode("utf-8"), add_bos=_add_bos, special=_special) if prompt != "" \
                else [self._vocab.bos_id]
            prompt_text = prompt
        else:
            tokens = [int(t) for t in prompt]
            prompt_text = None
        sampling = self._sampling(temperature, top_p, top_k, min_p, repeat_penalty, frequency_penalty, presence_penalty,
                                  typical_p, tfs_z, mirostat_mode, logit_bias, grammar, logits_processor)
        stops = [] if stop is None else ([stop] if isinstance(stop, str) else list(stop))
        stop_b = [s.encode("utf-8") for s in stops if s]
        cid = f"cmpl-{uuid.uuid4()}"
        created = int(time.time())
        model_name = model if model is not None else self.model_path
        if stream:
            return self._stream_completion(tokens, max_tokens, sampling, stop_b, seed, cid, created, model_name, cancel_event)
        req = self._run(tokens, max_tokens, sampling, stop_b, seed, cancel_event=cancel_event)
        text = req.text.decode("utf-8", errors="ignore")
        if echo and prompt_text is not None:
            text = prompt_text + text
        return {
            "id": cid, "object": "text_completion", "created": created, "model": model_name,
            "choices": [{"text": text, "index": 0, "logprobs": None, "finish_reason": req.finish_reason}],
            "usage": {"prompt_tokens": len(req.prompt), "completion_tokens": len(req.completion_tokens),
                      "total_tokens": len(req.prompt) + len(req.completion_tokens)},
        }

    def _stream_completion(self, tokens, max_tokens, sampling, stop_b, seed, cid, created, model_name,
                           cancel_event=None) -> Iterator[dict]:
        """Incremental streaming (upstream `_create_completion(stream=True)`): a chunk goes out as soon as its text can no
        longer change -- bytes that could still turn out to be the start of a stop string, or that end in the middle of a
        UTF-8 character, are held back. Closing the generator early cancels the generation."""
        q: "queue.Queue[Optional[int]]" = queue.Queue()
        cancel_event = cancel_event or threading.Event()
        req = self._make_request(tokens, max_tokens, sampling, stop_b, seed, on_token=q.put, cancel_event=cancel_event)
        self._scheduler.submit(req)

        def chunk(text, finish=None):
            return {"id": cid, "object": "text_completion", "created": created, "model": model_name,
                    "choices": [{"text": text, "index": 0, "logprobs": None, "finish_reason": finish}]}

        toks: List[int] = []
        sent = 0            # bytes of the completion already emitted
        try:
            while not req.done.is_set() or not q.empty():
                try:
                    t = q.get(timeout=0.05)
                except queue.Empty:
                    continue
                toks.append(t)
                raw = self._vocab.detokenize(toks)
                safe = len(raw)
                for sstr in stop_b:                     # nothing from a stop string on, and hold back a possible stop-string prefix
                    at = raw.find(sstr)
                    if at >= 0:
                        safe = min(safe, at)
                        continue
                    for i in range(min(len(sstr) - 1, len(raw)), 0, -1):
                        if raw.endswith(sstr[:i]):
                            safe = min(safe, len(raw) - i)
                            break
                while safe > sent:                      # hold back an incomplete UTF-8 character
                    try:
                        raw[sent:safe].decode("utf-8")
                        break
                    except UnicodeDecodeError as exc:
                        if exc.end < safe - sent and exc.reason != "unexpected end of data":
                            break                       # invalid bytes in the middle: emit with replacement below
                        safe = sent + exc.start
                if safe > sent:
                    yield chunk(raw[sent:safe].decode("utf-8", errors="ignore"))
                    sent = safe
            if req.error is not None:
                raise req.error
            tail = req.text[sent:] if len(req.text) > sent else b""
            if tail:
                yield chunk(tail.decode("utf-8", errors="ignore"))
            yield chunk("", req.finish_reason)
        finally:
            if not req.done.is_set():                   # the consumer went away: stop generating for it
                self._scheduler.cancel(req)

    __call__ = create_completion

    def create_chat_completion(self, messages: List[Dict[str, Any]], functions=None, function_call=None, tools=None,
                               tool_choice=None, temperature: float = 0.2, top_p: float = 0.95, top_k: int = 40,
                               min_p: float = 0.05, typical_p: float = 1.0, stream: bool = False,
                               stop: Optional[Union[str, List[str]]] = [], seed: Optional[int] = None,
                               response_format=None, max_tokens: Optional[int] = None, presence_penalty: float = 0.0,
                               frequency_penalty: float = 0.0, repeat_penalty: float = 1.1, tfs_z: float = 1.0,
                               mirostat_mode: int = 0, mirostat_tau: float = 5.0, mirostat_eta: float = 0.1,
                               model: Optional[str] = None, logits_processor=None, grammar=None, logit_bias=None,
                               logprobs: Optional[bool] = None, top_logprobs: Optional[int] = None,
                               cancel_event: Optional[threading.Event] = None, **kwargs: Any):
        """OpenAI-shaped chat completion; `messages` is not modified (the caller already owns and
        mutates it: api.py:39,44). `cancel_event` (not in upstream) abandons the generation when set."""
        if functions or tools or response_format is not None:
            raise NotImplementedError("functions / tools / response_format are not implemented")
        fmt = self._chat_formatter if self._chat_formatter is not None else cf.FORMATS[self.chat_format]
        res = fmt(messages=[dict(m) for m in messages])
        prompt = self.tokenize(res.prompt.encode("utf-8"), add_bos=not res.added_special, special=True)
        stops = [] if stop is None else ([stop] if isinstance(stop, str) else list(stop))
        if res.stop is not None:
            stops = stops + ([res.stop] if isinstance(res.stop, str) else list(res.stop))
        out = self.create_completion(prompt=prompt, temperature=temperature, top_p=top_p, top_k=top_k, min_p=min_p,
                                     typical_p=typical_p, stream=stream, stop=stops, seed=seed, max_tokens=max_tokens,
                                     presence_penalty=presence_penalty, frequency_penalty=frequency_penalty,
                                     repeat_penalty=repeat_penalty, tfs_z=tfs_z, mirostat_mode=mirostat_mode,
                                     mirostat_tau=mirostat_tau, mirostat_eta=mirostat_eta, model=model,
                                     logits_processor=logits_processor, grammar=grammar, logit_bias=logit_bias,
                                     cancel_event=cancel_event)
        if stream:
            return self._chat_chunks(out)
        return {
            "id": "chat" + out["id"], "object": "chat.completion", "created": out["created"], "model": out["model"],
            "choices": [{"index": 0, "message": {"role": "assistant", "content": out["choices"][0]["text"]},
                         "logprobs": None, "finish_reason": out["choices"][0]["finish_reason"]}],
            "usage": out["usage"],
        }

    @staticmethod
    def _chat_chunks(chunks: Iterator[dict]) -> Iterator[dict]:
        first = True
        for c in chunks:
            base = {"id": "chat" + c["id"], "object": "chat.completion.chunk", "created": c["created"], "model": c["model"]}
            if first:
                yield {**base, "choices": [{"index": 0, "delta": {"role": "assistant"}, "logprobs": None, "finish_reason": None}]}
                first = False
            ch = c["choices"][0]
            delta = {"content": ch["text"]} if ch["finish_reason"] is None else {}
            yield {**base, "choices": [{"index": 0, "delta": delta, "logprobs": None, "finish_reason": ch["finish_reason"]}]}
