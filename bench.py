#!/usr/bin/env python
"""Headline benchmark: decode tok/s of the Llama-3-8B-arch Q4_K_M synthetic GGUF at batch 1
(BASELINE.json configs[1]) on N B200s (N independent replicas, weak scaling).

    python bench.py --gpus 1 --steps 256 --warmup 16
    torchrun ... bench.py --gpus N ...            (one rank per GPU)
    python bench.py --impl reference              (the CPU restatement of the reference's
                                                   llama.cpp CPU path, on the host cores)

One "step" = one decoded token for one sequence (every weight byte and every live KV byte read
once). `value` times K steps with CUDA events on the engine's stream with the sequence state
resident in HBM; `e2e` times the same K tokens through the public `Llama.create_completion`
call (host token ids in, host text out). See DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (arch, ftype, prompt tokens, sequences per step)   -- BASELINE.json configs
    "llama3-8b-q4_k_m-bs1": ("llama-3-8b", "Q4_K_M", 512, 1),
    "tinyllama-1.1b-q4_0-bs1": ("tinyllama-1.1b", "Q4_0", 128, 1),
    "llama3-8b-q8_0-bs1": ("llama-3-8b", "Q8_0", 512, 1),
    "llama3-8b-q8_0-bs32": ("llama-3-8b", "Q8_0", 512, 32),
    "llama3-8b-q4_k_m-bs32": ("llama-3-8b", "Q4_K_M", 512, 32),
    "mistral-7b-q4_k_m-8k": ("mistral-7b", "Q4_K_M", 8192 - 512, 1),
    "llama3-70b-q4_k_m-tp": ("llama-3-70b", "Q4_K_M", 128, 1),       # configs[4]: tensor parallel over all ranks (N = 1: one GPU)
    "tiny-q4_k_m-tp": ("tiny", "Q4_K_M", 32, 1),
    "tiny-q4_k_m": ("tiny", "Q4_K_M", 32, 1),
    "tiny-q4_k_m-bs8": ("tiny", "Q4_K_M", 32, 8),
}


def batched_decode_run(path, n_prompt, bs, n_warm, n_steps, device, seed=7):
    """bs sequences with n_prompt-token prompts resident in the paged KV cache, then n_steps decode steps of all of
    them at once (b2_decode): CUDA events on the engine stream around the timed steps. Also returns the p50 of the
    per-prompt prefill times (host-timed synchronous b2_prefill calls) seen while filling the slots."""
    import numpy as np
    from llama_cpp.engine import Engine, SamplingSettings
    from llama_cpp.gguf_file import GGUFFile
    eng = Engine(GGUFFile(path), n_ctx=n_prompt + n_warm + n_steps + 16, max_seqs=bs, device=device, kv_block_tokens=64)
    greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)
    rng = np.random.default_rng(seed)
    pre = []
    with eng.lock:
        for s in range(bs):
            prompt = [int(t) for t in rng.integers(0, eng.dims.n_vocab, size=n_prompt)]
            eng.ensure_pages(s, n_prompt + n_warm + n_steps + 8)
            eng.seq_begin(s, 0, [], greedy)
            t0 = time.perf_counter()
            eng.prefill(s, prompt)
            pre.append((time.perf_counter() - t0) * 1e3)
            eng.sample(s)
        slots = list(range(bs))
        for _ in range(n_warm):
            eng.decode(slots)
        l0 = eng.launch_count()
        eng.timer_start()
        for _ in range(n_steps):
            eng.decode(slots)
        ms = eng.timer_stop()
        launches = eng.launch_count() - l0
    eng.close()
    pre = sorted(pre[1:] or pre)            # the first call also allocates the prompt workspace
    return dict(ms=ms, launches=launches, prefill_p50_ms=pre[len(pre) // 2], prefill_n=len(pre))


def http_burst(path, batch, rank, barrier, n_ctx=1024, port0=18200):
    """`batch` concurrent POST /response against an in-process uvicorn server (llama_cpp.server) that fronts one engine
    replica on this rank's GPU. Returns generated tokens, wall time of the burst and what the scheduler saw."""
    import asyncio
    import aiohttp
    import numpy as np
    import uvicorn
    from llama_cpp import Llama, server
    llm = Llama(model_path=path, n_gpu_layers=-1, n_ctx=n_ctx, seed=1, verbose=False, n_seq_max=batch)
    app = server.create_app([llm], max_queue_size=2 * batch, permits=batch, consumers=batch, timeout_seconds=600.0,
                            max_context_tokens=n_ctx)
    port = port0 + rank
    srv = uvicorn.Server(uvicorn.Config(app, host="127.0.0.1", port=port, log_level="warning"))
    th = threading.Thread(target=srv.run, daemon=True)
    th.start()
    for _ in range(200):
        if srv.started:
            break
        time.sleep(0.05)
    rng = np.random.default_rng(7 + rank)
    words = ["time", "river", "blue", "seven", "music", "open", "garden", "light", "story", "window", "coffee", "travel"]

    def body(i):
        ctx = [{"turn": "user" if j % 2 == 0 else "bot", "message": " ".join(rng.choice(words, size=12))} for j in range(7)]
        return {"bot_profile": {"name": f"Ann{i}.f", "appearance": "a,b,c,tall,green eyes", "system_prompt": ""},
                "user_profile": {"name": "u"}, "context": ctx}

    async def burst(bodies):
        async with aiohttp.ClientSession(timeout=aiohttp.ClientTimeout(total=900)) as sess:
            async def one(bd):
                async with sess.post(f"http://127.0.0.1:{port}/response", json=bd) as r:
                    return r.status, len((await r.json()).get("response", "")) if r.status == 200 else 0
            return await asyncio.gather(*[one(bd) for bd in bodies])

    asyncio.run(burst([body(-1)]))                        # warm the path (graphs, workspace)
    st0 = dict(llm._scheduler.stats)
    llm._scheduler.stats["first_submit"] = 0.0
    bodies = [body(i) for i in range(batch)]
    req_bytes = sum(len(json.dumps(bd)) for bd in bodies)
    barrier()
    t0 = time.perf_counter()
    res = asyncio.run(burst(bodies))
    wall = time.perf_counter() - t0
    barrier()
    st = llm._scheduler.stats
    n_gen = st["tokens"] - st0["tokens"] + batch          # + the first token of every request (sampled at admission)
    n_prompt_tok = st["prefill_tokens"] - st0["prefill_tokens"] + st["reused_tokens"] - st0["reused_tokens"]
    steps = max(1, st["steps"] - st0["steps"])
    out = {"value": n_gen / wall, "unit": "tok/s", "h2d_bytes_per_step": round(4.0 * n_prompt_tok / steps, 1),
           "d2h_bytes_per_step": round(4.0 * n_gen / steps, 1), "tokens": int(n_gen), "wall_s": wall,
           "http_status": sorted(set(c for c, _ in res)), "requests": batch, "request_bytes": req_bytes,
           "response_chars": int(sum(n for _, n in res)), "prompt_tokens": int(n_prompt_tok), "max_batch": st["max_batch"],
           "scheduler_s": {k: round(st[k] - st0[k], 3) for k in ("t_admit", "t_decode", "t_host")}, "decode_steps": int(steps),
           "timeline_s": {"first_submit": round(st["first_submit"] - t0, 3), "last_submit": round(st["last_submit"] - t0, 3),
                          "last_done": round(st["last_done"] - t0, 3)},
           "note": f"{batch} concurrent POST /response (chat bodies, the reference's sampling settings, generation to the "
                   f"{n_ctx}-token window) against llama_cpp.server over a loopback socket with permits = consumers = {batch}; "
                   "generated tokens / wall time of the burst: JSON parsing, prompt assembly, tokenisation, prompt passes, "
                   "batched decode steps, detokenisation and the HTTP responses are all inside"}
    srv.should_exit = True
    th.join(timeout=10)
    llm.close()
    return out


def model_path(arch, ftype, seed=1):
    base = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    return os.path.join(base, "b200_gguf_models", f"{arch}-{ftype}-seed{seed}.gguf")


def ensure_model(arch, ftype, seed=1):
    from llama_cpp import synth
    path = model_path(arch, ftype, seed)
    if not os.path.exists(path):
        synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=0.1, with_vocab=True)
    return path


def algorithmic_bytes(gguf, ctx_mean):
    """W_read (all tensors but token_embd, plus one embedding row) + ctx * KV bytes."""
    md = gguf.metadata
    total = sum(t.nbytes for t in gguf.tensors.values())
    emb = gguf.tensors["token_embd.weight"]
    w_read = total - emb.nbytes + emb.nbytes // emb.shape[1]
    n_head = int(md["llama.attention.head_count"])
    kv = int(md.get("llama.attention.head_count_kv", n_head))
    hd = int(md["llama.embedding_length"]) // n_head
    kv_pos = int(md["llama.block_count"]) * 2 * kv * hd * 2
    return w_read, kv_pos, w_read + ctx_mean * kv_pos


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic_bytes(workload):
    """dram read+write bytes per decode step from the committed ncu capture of THIS workload, else None (a figure
    captured on another model says nothing about this one)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t.get("dram_bytes_per_step") if t.get("workload") == workload else None
    except Exception:
        return None


def cpu_reference_run(path, n_warm, n_steps, budget_s, prompt_len=16):
    """Time the CPU restatement of the reference's llama.cpp CPU path: decode tokens one at a time."""
    # torchrun exports OMP_NUM_THREADS=1 to its workers and the OpenMP runtime may already be up (torch): the thread count
    # is set through the library's own entry point. The CPU arm is meant to use every host core it can.
    n_threads = int(os.environ.get("B2_CPU_THREADS", "0")) or (len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count())
    from oracle.cpu_ref import CpuRef
    import numpy as np
    c = CpuRef(path, n_ctx=prompt_len + n_warm + n_steps + 8, n_threads=n_threads)
    rng = np.random.default_rng(0)
    toks = rng.integers(0, c.n_vocab, size=prompt_len + n_warm + n_steps)
    t0 = time.perf_counter()
    c.eval(int(toks[0]))
    first = time.perf_counter() - t0
    for t in toks[1:prompt_len]:
        c.eval(int(t), want_logits=False)
    per = max(first, 1e-4)
    n_warm = max(0, min(n_warm, int(0.2 * budget_s / per)))
    n_steps = max(1, min(n_steps, int(0.8 * budget_s / per)))
    i = prompt_len
    for _ in range(n_warm):
        c.eval(int(toks[i])); i += 1
    t0 = time.perf_counter()
    for _ in range(n_steps):
        c.eval(int(toks[i])); i += 1
    dt = time.perf_counter() - t0
    threads, simd = c.threads, c.simd
    c.close()
    return dict(value=n_steps / dt, ms_per_step=dt / n_steps * 1e3, steps=n_steps, warmup=n_warm, cores=threads,
                sample=f"{n_steps} single-token decode steps (all layers + lm_head) after a {prompt_len}-token context, "
                       f"{simd} integer dots on {threads} OpenMP threads; KV reads are <2% of the bytes at the workload's "
                       f"context, weights are streamed in full")


def parity_check(eng, path, n_vocab, n_prompt=6, n_new=6, seed=5):
    """Run a short greedy generation on the TIMED engine instance (same weights in HBM, same CUDA graphs) and teacher-force
    the C restatement of the reference's llama graph (oracle/cpu_ref.c) over the same tokens: all layers at the benchmark
    geometry, checked inside the benchmark process. The restatement runs in its checker mode (float activations against
    dequantised weights, double accumulation): its timed mode quantises activations to int8 as ggml does, and on a
    32-layer random-init model that noise alone is ~10% of the logit spread (measured, round 2). Raises if it is off."""
    import numpy as np
    from llama_cpp.engine import SamplingSettings
    from oracle.cpu_ref import CpuRef
    rng = np.random.default_rng(seed)
    prompt = [int(t) for t in rng.integers(0, n_vocab, size=n_prompt)]
    slot = 0
    with eng.lock:
        eng.pages.release(slot)
        eng.ensure_pages(slot, n_prompt + n_new + 2)
        eng.seq_begin(slot, 0, [], SamplingSettings(temperature=0.0, repeat_penalty=1.0))
        eng.prefill(slot, prompt)                 # < 8 tokens: through the mat-vec kernels, like the timed decode steps
        gpu_logits, toks = [eng.logits(slot)], [eng.sample(slot)]
        for _ in range(n_new - 1):
            toks.append(eng.decode([slot])[0])
            gpu_logits.append(eng.logits(slot))
    c = CpuRef(path, n_ctx=n_prompt + n_new + 4, n_threads=len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else 0)
    c.set_exact(True)
    try:
        for t in prompt[:-1]:
            c.eval(t, want_logits=False)
        seq = [prompt[-1]] + toks[:-1]
        worst, agree, clear, std = 0.0, 0, 0, 0.0
        for t, g in zip(seq, gpu_logits):
            r = c.eval(int(t))
            worst = max(worst, float(np.abs(g - r).max()))
            std = max(std, float(r.std()))
            top2 = np.partition(r, -2)[-2:]
            if top2[1] - top2[0] > 0.05:
                clear += 1
                agree += int(np.argmax(r) == np.argmax(g))
    finally:
        c.set_exact(False)
        c.close()
    # One layer: <= 5e-3 of the logit std (tests/test_gpu_wide.py). The differences (f16 KV entries that round one ulp apart,
    # summation order) compound roughly linearly through the layers of a random-init model: measured 1.3e-2 on the 32-layer
    # 8B and 2.8e-2 on the 80-layer 70B, correlation 1.00000 and every clear arg-max equal. Bar: (5e-3 + 5e-4 per layer).
    n_layer = eng.dims.n_layer
    bar = 5e-3 + 5e-4 * n_layer
    ok = worst <= bar * std and agree == clear
    out = {"max_abs_logit_diff": round(worst, 4), "logit_std": round(std, 3), "frac_of_logit_std": round(worst / max(std, 1e-9), 5),
           "bar_frac_of_logit_std": round(bar, 4), "argmax_agree": f"{agree}/{clear}", "steps": n_new, "ok": bool(ok),
           "against": "oracle/cpu_ref.c in checker mode (float activations x dequantised weights, double accumulation), "
                      "teacher-forced on the engine's greedy tokens; every layer of the timed model"}
    if not ok:
        raise RuntimeError(f"parity check failed: {out}")
    return out


def tp_workload(args, cfg, path, n_prompt, rank, world, local_rank, barrier, torch, dist):
    """BASELINE.json configs[4]: ONE model sharded over all ranks (column / row splits of the quantised weights, NCCL
    all-reduce after attn_output and ffn_down, all-gather of the logits); batch-1 greedy decode. Strong scaling: the work
    is one sequence whatever N is. N = 1 runs the whole model on one GPU (the 70B Q4_K_M file is 42.5 GB)."""
    import numpy as np
    from llama_cpp import Llama, replicas
    from llama_cpp.engine import Engine, SamplingSettings
    from llama_cpp.gguf_file import GGUFFile
    K, W = args.steps, args.warmup
    tp = None
    if world > 1:
        ids = [Engine.tp_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        tp = (rank, world, ids[0])
    n_ctx = n_prompt + W + 2 * K + 32
    t_load = time.perf_counter()
    llm = Llama(model_path=path, n_gpu_layers=-1, n_ctx=n_ctx, seed=1, verbose=False, tp=tp)
    t_load = time.perf_counter() - t_load
    eng = llm._engine
    V = eng.dims.n_vocab
    prompt = [int(t) for t in np.random.default_rng(123).integers(0, V, size=n_prompt)]      # the same on every rank
    greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)
    with eng.lock:
        eng.ensure_pages(0, n_ctx)
        eng.seq_begin(0, 0, [], greedy)
        t0 = time.perf_counter()
        eng.prefill(0, prompt)
        prefill_ms = (time.perf_counter() - t0) * 1e3
        first = eng.sample(0)
        warm = eng.decode_n(0, W)
        clocks = ClockSampler(local_rank)
        clocks.start()
        launches0 = eng.launch_count()
        barrier()
        eng.timer_start()
        toks = eng.decode_n(0, K)
        ms = eng.timer_stop()
        barrier()
        launches = eng.launch_count() - launches0
        clk = clocks.result()
        kinds = eng.profile_decode_step(0)
        mode = eng.mode()
    ms_max = replicas.dist_max([ms])[0]
    if world > 1:       # every rank must have produced the same tokens
        allt = [None] * world
        dist.all_gather_object(allt, [first] + warm + toks)
        if any(t != allt[0] for t in allt):
            raise RuntimeError("tensor-parallel ranks disagree on the generated tokens")
    value = K / (ms_max / 1e3)
    # end to end through the public API: host ids in, text out, every rank makes the same call
    llm.reset()
    barrier()
    t0 = time.perf_counter()
    out = llm.create_completion(prompt, max_tokens=K, temperature=0.0, repeat_penalty=1.0)
    wall = replicas.dist_max([time.perf_counter() - t0])[0]
    barrier()
    n_gen = out["usage"]["completion_tokens"]
    e2e = {"value": n_gen / wall, "unit": "tok/s", "h2d_bytes_per_step": round(4 * n_prompt / max(n_gen, 1), 2), "d2h_bytes_per_step": 4,
           "tokens": n_gen, "wall_s": wall, "ttft_ms": llm.last_timings["ttft_ms"],
           "note": "whole Llama.create_completion(token ids, max_tokens=K, temperature=0) call on every rank at once (prompt token "
                   "by token through the mat-vec path in tensor-parallel mode), host-timed, max over ranks"}
    if rank != 0:
        llm.close()
        return 0
    g = GGUFFile(path)
    ctx_mean = n_prompt + W + K / 2.0
    w_read, kv_pos, step_bytes = algorithmic_bytes(g, ctx_mean)
    peak, peak_src = measured_peak_gbs()
    achieved = step_bytes / (ms_max / K / 1e3) / 1e9
    n_coll = 2 * eng.dims.n_layer + 1 if world > 1 else 0
    roof = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak * world, "unit": "GB/s", "frac": round(achieved / (peak * world), 4),
            "traffic": None, "peak_source": peak_src + f" x {world} GPUs", "bytes_per_launch": int(step_bytes),
            "launch": "one decode step of the sharded model = one CUDA-graph launch per rank; algorithmic bytes = W_read + mean_ctx * "
                      "KV bytes/position over all ranks",
            "kernels": {k: {"ms": round(v["ms"], 4), "launches": v["launches"]} for k, v in kinds.items()},
            "kernels_note": "rank 0, one un-graphed step with events around every launch; attn_out and ffn_down include their all-reduce, "
                            "lm_head its all-gather",
            "collectives_per_step": n_coll,
            "limiting_collective": (f"all-reduce of {eng.dims.n_embd} floats ({eng.dims.n_embd * 4 // 1024} KB) after attn_output and ffn_down: "
                                    f"{2 * eng.dims.n_layer} per token, latency-bound") if world > 1 else None}
    line = {"metric": "decode_tokens_per_s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_max / K, 5), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int8x2 activations x int4/5/6 weights (dp4a), f32 accumulate and all-reduce, f16 KV", "data": "synthetic",
            "config": cfg, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "engine": mode,
            "prefill_ms": round(prefill_ms, 2), "load_s": round(t_load, 2), "w_read_bytes": int(w_read), "kv_bytes_per_pos": kv_pos,
            "ctx_mean": ctx_mean}
    if not args.no_cpu_baseline:
        try:
            r = cpu_reference_run(path, 1, 8, budget_s=args.cpu_budget_s)
            line["cpu_baseline"] = {"value": round(r["value"], 3), "unit": "tok/s", "cores": r["cores"], "kind": "port", "sample": r["sample"]}
        except Exception as exc:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {exc}"}
    if world == 1 and not args.no_parity_check:
        line["parity_check"] = parity_check(eng, path, V, n_new=3)
    print(json.dumps(line))
    llm.close()
    return 0


def batch_workload(args, cfg, path, n_prompt, batch, rank, world, local_rank, barrier, torch, dist):
    """Continuous-batching configs: `batch` sequences decode together, one weight pass per step (tensor-core GEMM path)."""
    import numpy as np
    from llama_cpp import Llama
    from llama_cpp.gguf_file import GGUFFile
    K, W = args.steps, args.warmup
    clocks = ClockSampler(local_rank)
    if not args.ncu_mode:
        clocks.start()
    barrier()
    r = batched_decode_run(path, n_prompt, batch, W, K, local_rank)
    barrier()
    clk = clocks.result() if not args.ncu_mode else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    from llama_cpp import replicas
    agg = replicas.aggregate_throughput(batch * K, r["ms"])       # all ranks' tokens / the slowest rank's device time
    ms_max, value = agg["ms"], agg["value"]
    e2e = None
    if not args.ncu_mode:
        # end to end: `batch` concurrent callers of the public API (as /response handlers would be), host ids in, text out
        os.environ["B200_MAX_SEQS"] = str(batch)
        n_new = min(K, 64)
        # through the HTTP API, as BASELINE.json words it ("continuous batching bs32 via /response"): this rank's replica
        # behind llama_cpp.server (the reference's app with permits / consumers / queue raised from 1 / 1 / 5 to the
        # batch size), `batch` concurrent POST /response over a real socket
        e2e = http_burst(path, batch, rank, barrier)
        agg2 = replicas.aggregate_throughput(e2e["tokens"], e2e["wall_s"] * 1e3)
        e2e["value"] = agg2["value"]
    if rank != 0:
        return 0
    g = GGUFFile(path)
    ctx_mean = n_prompt + W + K / 2.0
    w_read, kv_pos, _ = algorithmic_bytes(g, ctx_mean)
    step_bytes = w_read + batch * ctx_mean * kv_pos
    peak, peak_src = measured_peak_gbs()
    achieved = step_bytes / (ms_max / K / 1e3) / 1e9
    roof = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
            "traffic": None, "peak_source": peak_src, "bytes_per_launch": int(step_bytes),
            "launch": f"one batched decode step ({batch} sequences): weights once + every sequence's KV; algorithmic bytes = "
                      "W_read + batch * mean_ctx * KV bytes/position"}
    line = {"metric": "decode_tokens_per_s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_max / K, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 activations x int4/6/8 weights expanded to f16 (tcgen05 kind::f16), f32 accumulate, f16 KV",
            "data": "synthetic", "config": cfg, "clocks": clk, "e2e": e2e, "gpu_launches": int(r["launches"]), "roofline": roof,
            "prefill_p50_ms": round(r["prefill_p50_ms"], 2), "prefill_tok_s": round(n_prompt / r["prefill_p50_ms"] * 1e3, 1),
            "w_read_bytes": int(w_read), "kv_bytes_per_pos": kv_pos, "ctx_mean": ctx_mean}
    if not args.no_cpu_baseline and not args.ncu_mode:
        try:
            c = cpu_reference_run(path, 1, 64, budget_s=args.cpu_budget_s)
            line["cpu_baseline"] = {"value": round(c["value"], 3), "unit": "tok/s", "cores": c["cores"], "kind": "port",
                                    "sample": c["sample"] + " (one sequence; the CPU path has no batching)"}
        except Exception as exc:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {exc}"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama3-8b-q4_k_m-bs1", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--batch-extra", type=int, default=32, help="bs1 workloads: also time this many sequences per step (0 = skip)")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    ap.add_argument("--profile-kernels", action="store_true", help="also print a per-kernel-kind breakdown to stderr")
    ap.add_argument("--ncu-mode", action="store_true", help="short run for ncu: skip e2e, cpu baseline, clocks")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    arch, ftype, n_prompt, batch = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = {"workload": f"{args.workload}: {arch}-arch {ftype} synthetic GGUF (seed 1), {n_prompt}-token prompt then "
                       f"batch-{batch} greedy decode", "prompt_tokens": n_prompt, "batch": batch,
           "parallelism": f"replicas x{world}" if world > 1 else "single GPU",
           "l2": "every step streams all weights (0.6-8 GB, larger than the 126 MB L2); no flush needed"}

    if args.impl == "reference":
        # the reference's own CPU implementation of the path (restated: llama_cpp is not installable here)
        if rank != 0:
            return 0
        path = ensure_model(arch, ftype)
        r = cpu_reference_run(path, args.warmup, args.steps, budget_s=150.0)
        line = {"impl": "reference", "metric": "decode_tokens_per_s", "value": r["value"], "unit": "tok/s", "n_gpus": 0,
                "steps": r["steps"], "steps_requested": args.steps, "warmup": r["warmup"], "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 x int4/6 dot, f32 accumulate",
                "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": r["value"], "unit": "tok/s", "cores": r["cores"], "kind": "port", "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    os.environ["B200_DEVICE"] = str(local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if local_rank == 0:
        path = ensure_model(arch, ftype)
    barrier()
    path = model_path(arch, ftype)

    from llama_cpp import Llama
    from llama_cpp.engine import SamplingSettings
    from llama_cpp.gguf_file import GGUFFile
    K, W = args.steps, args.warmup
    if args.workload.endswith("-tp"):
        cfg["parallelism"] = f"tensor parallel x{world} (column/row weight splits, NCCL all-reduce)" if world > 1 else "single GPU"
        rc = tp_workload(args, cfg, path, n_prompt, rank, world, local_rank, barrier, torch, dist)
        if world > 1:
            dist.destroy_process_group()
        return rc
    if batch > 1:
        return batch_workload(args, cfg, path, n_prompt, batch, rank, world, local_rank, barrier, torch, dist)
    n_ctx = n_prompt + W + K + 16
    t_load = time.perf_counter()
    llm = Llama(model_path=path, n_gpu_layers=-1, n_ctx=n_ctx, seed=1, verbose=False)
    t_load = time.perf_counter() - t_load
    eng = llm._engine
    rng = np.random.default_rng(123 + rank)
    prompt = [int(t) for t in rng.integers(0, eng.dims.n_vocab, size=n_prompt)]

    # ---- kernel-resident timing: prompt KV and all state already in HBM -------------------------
    greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)
    with eng.lock:
        eng.ensure_pages(0, n_ctx)
        eng.seq_begin(0, 0, [], greedy)
        t0 = time.perf_counter()
        eng.prefill(0, prompt)
        ttft_cold_ms = (time.perf_counter() - t0) * 1e3     # includes the one-time prompt workspace allocation
        eng.seq_begin(0, 0, [], greedy)
        t0 = time.perf_counter()
        eng.prefill(0, prompt)                              # same prompt again, warm: host-timed, synchronous call
        ttft_ms = (time.perf_counter() - t0) * 1e3
        eng.sample(0)
        eng.decode_n(0, W)                         # warm-up (also captures the CUDA graph)
        clocks = ClockSampler(local_rank)
        if not args.ncu_mode:
            clocks.start()
        launches0 = eng.launch_count()
        barrier()
        eng.timer_start()
        eng.decode_n(0, K)
        ms = eng.timer_stop()
        barrier()
        launches = eng.launch_count() - launches0
        clk = clocks.result() if not args.ncu_mode else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        kinds = eng.profile_decode_step(0) if (args.profile_kernels or not args.ncu_mode) else None
        mode = eng.mode()
    from llama_cpp import replicas
    agg = replicas.aggregate_throughput(K, ms)                    # all ranks' tokens / the slowest rank's device time
    ms_max, value = agg["ms"], agg["value"]

    # ---- end to end through the public API: host token ids in, host text out -------------------
    e2e = None
    if not args.ncu_mode:
        llm.reset()
        prompt2 = [int(x) for x in rng.integers(0, eng.dims.n_vocab, size=n_prompt)]
        llm.create_completion(prompt2[:64], max_tokens=8, temperature=0.0, repeat_penalty=1.0)     # warm the API path
        barrier()
        t0 = time.perf_counter()
        out = llm.create_completion(prompt2, max_tokens=K, temperature=0.0, repeat_penalty=1.0)
        wall = time.perf_counter() - t0
        barrier()
        tm = llm.last_timings
        n_gen = out["usage"]["completion_tokens"]
        t2 = replicas.dist_max([wall, tm["ttft_ms"], tm["decode_ms"] / 1e3])
        n_gen_all = replicas.dist_sum([float(n_gen)])[0]
        e2e = {"value": n_gen_all / t2[0], "unit": "tok/s",
               "h2d_bytes_per_step": round(4 * n_prompt / max(n_gen, 1), 2), "d2h_bytes_per_step": 4,
               "tokens": n_gen, "ttft_ms": t2[1], "wall_s": wall,
               "decode_phase_tok_s": world * max(n_gen - 1, 1) / t2[2],
               "note": "WHOLE Llama.create_completion(token ids, max_tokens=K, temperature=0) call, host-timed: H2D of the "
                       f"{n_prompt} prompt ids, the prompt pass, K decode steps with per-chunk D2H of the sampled ids, stop "
                       "checks, detokenise; generated tokens / wall time (decode_phase_tok_s = first token to last only)"}

    # ---- the other headline numbers of BASELINE.json on the same model: prompt p50 and 32 sequences per step
    extra = None
    if not args.ncu_mode and args.batch_extra > 0:
        try:
            bsteps = min(K, 64)
            br = batched_decode_run(path, n_prompt, args.batch_extra, 4, bsteps, local_rank)
            extra = {"bs": args.batch_extra, "value": round(args.batch_extra * bsteps / (br["ms"] / 1e3), 1), "unit": "tok/s",
                     "ms_per_step": round(br["ms"] / bsteps, 4), "steps": bsteps, "prompt_tokens": n_prompt,
                     "prefill_p50_ms": round(br["prefill_p50_ms"], 2), "prefill_runs": br["prefill_n"],
                     "prefill_tok_s": round(n_prompt / br["prefill_p50_ms"] * 1e3, 1),
                     "note": "per GPU; b2_decode over all slots per step (tensor-core GEMM path), CUDA events on the engine "
                             "stream; prefill p50 = synchronous b2_prefill calls, host-timed"}
        except Exception as exc:  # noqa: BLE001
            extra = {"error": str(exc)}
    if rank != 0:
        return 0
    g = GGUFFile(path)
    ctx_mean = n_prompt + W + K / 2.0
    w_read, kv_pos, step_bytes = algorithmic_bytes(g, ctx_mean)
    peak, peak_src = measured_peak_gbs()
    achieved = step_bytes / (ms_max / K / 1e3) / 1e9
    roof = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
            "traffic": ncu_traffic_bytes(args.workload), "peak_source": peak_src, "bytes_per_launch": int(step_bytes),
            "launch": "one decode step = one CUDA-graph launch (embed, 32 x [qkv, attention, attn_out, gate_up, ffn_down], "
                      "lm_head, sampler); algorithmic bytes = W_read + mean_ctx * KV bytes/position",
            "frac_of_8TBs_nominal": round(achieved / 8000.0, 4)}
    if kinds:
        roof["kernels"] = {k: {"ms": round(v["ms"], 4), "launches": v["launches"],
                               "GBps": round(v["bytes"] / max(v["ms"], 1e-9) / 1e6, 1)} for k, v in kinds.items()}
        roof["kernels_note"] = "one un-graphed decode step with events around every launch (no overlap between kernels)"
        top = max((k for k in kinds if kinds[k]["bytes"] > 0), key=lambda k: kinds[k]["ms"])
        kv = kinds[top]
        roof["dominant_kernel"] = {"name": f"matvec_kernel, {top} launches" if top != "attention" else "attn_decode_kernel",
                                   "launches_per_step": kv["launches"], "bytes_per_launch": int(kv["bytes"] / kv["launches"]),
                                   "us_per_launch": round(kv["ms"] / kv["launches"] * 1e3, 2),
                                   "achieved": round(kv["bytes"] / max(kv["ms"], 1e-9) / 1e6, 1), "unit": "GB/s",
                                   "frac": round(kv["bytes"] / max(kv["ms"], 1e-9) / 1e6 / peak, 4),
                                   "note": "CUDA events around each launch of this kind in one un-graphed step (cold start of every "
                                           "launch included); the whole-step figure above is the graph replay"}
    line = {"metric": "decode_tokens_per_s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_max / K, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8x2 activations x int4/6 weights (dp4a), f32 accumulate, f16 KV", "data": "synthetic",
            "config": cfg, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof,
            "engine": mode, "prefill_ms": round(ttft_ms, 2), "prefill_tok_s": round(n_prompt / ttft_ms * 1e3, 1),
            "prefill_cold_ms": round(ttft_cold_ms, 2), "load_s": round(t_load, 2),
            "w_read_bytes": int(w_read), "kv_bytes_per_pos": kv_pos, "ctx_mean": ctx_mean, "batched": extra}
    if not args.no_cpu_baseline and not args.ncu_mode:
        try:
            r = cpu_reference_run(path, 1, 64, budget_s=args.cpu_budget_s)
            line["cpu_baseline"] = {"value": round(r["value"], 3), "unit": "tok/s", "cores": r["cores"], "kind": "port",
                                    "sample": r["sample"]}
        except Exception as exc:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {exc}"}
    if not args.ncu_mode and not args.no_parity_check:
        line["parity_check"] = parity_check(eng, path, eng.dims.n_vocab)      # raises (no JSON line) if the timed engine is off
    print(json.dumps(line))
    llm.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
