"""Time the prompt path (b2_prefill) on a synthetic model: warm, host-timed, synchronous calls.
usage: python scripts/prefill_time.py [arch] [ftype] [n_prompt ...]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llama-fastapi-k8s-gpu_b200"))
import numpy as np
import bench
from llama_cpp.engine import Engine, SamplingSettings
from llama_cpp.gguf_file import GGUFFile

arch = sys.argv[1] if len(sys.argv) > 1 else "llama-3-8b"
ftype = sys.argv[2] if len(sys.argv) > 2 else "Q4_K_M"
lens = [int(a) for a in sys.argv[3:]] or [512]
path = bench.ensure_model(arch, ftype)
eng = Engine(GGUFFile(path), n_ctx=max(lens) + 64, max_seqs=1, kv_block_tokens=64)
greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)
rng = np.random.default_rng(1)
eng.ensure_pages(0, max(lens) + 32)
for n in lens:
    prompt = [int(t) for t in rng.integers(0, eng.dims.n_vocab, size=n)]
    best = 1e9
    for rep in range(int(os.environ.get("REPS", "3"))):
        eng.seq_begin(0, 0, [], greedy)
        t0 = time.perf_counter()
        eng.prefill(0, prompt)
        best = min(best, (time.perf_counter() - t0) * 1e3)
    print(f"prefill n={n}: {best:.2f} ms  {n / best * 1e3:.0f} tok/s", flush=True)
eng.close()
