"""Time batched decode (b2_decode with bs slots per step): host-timed over synchronous calls.
usage: python scripts/batch_time.py [arch] [ftype] [prompt_len] [bs ...]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llama-fastapi-k8s-gpu_b200"))
import numpy as np
import bench
from llama_cpp.engine import Engine, SamplingSettings
from llama_cpp.gguf_file import GGUFFile

arch = sys.argv[1] if len(sys.argv) > 1 else "llama-3-8b"
ftype = sys.argv[2] if len(sys.argv) > 2 else "Q8_0"
n_prompt = int(sys.argv[3]) if len(sys.argv) > 3 else 512
sizes = [int(a) for a in sys.argv[4:]] or [32]
steps = int(os.environ.get("STEPS", "32"))
path = bench.ensure_model(arch, ftype)
eng = Engine(GGUFFile(path), n_ctx=n_prompt + steps + 80, max_seqs=max(sizes), kv_block_tokens=64)
greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)
rng = np.random.default_rng(1)
for bs in sizes:
    for s in range(bs):
        prompt = [int(t) for t in rng.integers(0, eng.dims.n_vocab, size=n_prompt)]
        eng.ensure_pages(s, n_prompt + steps + 40)
        eng.seq_begin(s, 0, [], greedy)
        eng.prefill(s, prompt)
        eng.sample(s)
    slots = list(range(bs))
    for _ in range(4):
        eng.decode(slots)
    t0 = time.perf_counter()
    for _ in range(steps):
        eng.decode(slots)
    dt = time.perf_counter() - t0
    print(f"bs={bs}: {dt / steps * 1e3:.3f} ms/step  {bs * steps / dt:.0f} tok/s", flush=True)
    for s in range(bs):
        eng.release(s)
eng.close()
