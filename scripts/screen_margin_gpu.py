"""Screen prompt seeds ON THE GPU for the designated parity run (BASELINE.json configs[0]): for one synthetic model,
greedy-decode n_new tokens from many random prompts with the engine and report the prompts whose smallest top-2 logit
margin over the run is largest. The winners are then re-checked on the CPU with the float32 oracle
(tests/test_gpu_cfg1_parity.py records the chosen seed and the oracle's margin).
usage: python scripts/screen_margin_gpu.py arch ftype model_seed output_std n_prompt_seeds [n_prompt n_new]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")]
from llama_cpp import synth
from llama_cpp.engine import Engine, SamplingSettings
from llama_cpp.gguf_file import GGUFFile

arch, ftype, mseed, ostd, n_seeds = sys.argv[1], sys.argv[2], int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5])
n_prompt = int(sys.argv[6]) if len(sys.argv) > 6 else 128
n_new = int(sys.argv[7]) if len(sys.argv) > 7 else 128
path = f"/tmp/screen_{arch}_{ftype}_{mseed}.gguf"
synth.write_synthetic_model(path, arch, ftype, seed=mseed, output_std=ostd, with_vocab=False)
eng = Engine(GGUFFile(path), n_ctx=n_prompt + n_new + 16, max_seqs=1)
V = eng.dims.n_vocab
res = []
t0 = time.time()
for ps in range(n_seeds):
    prompt = [int(t) for t in np.random.default_rng(1000 + ps).integers(0, V, n_prompt)]
    eng.pages.release(0)
    eng.ensure_pages(0, n_prompt + n_new + 1)
    eng.seq_begin(0, 0, [], SamplingSettings(temperature=0.0, repeat_penalty=1.0))
    eng.prefill(0, prompt)
    margins = []
    tok = None
    for i in range(n_new):
        if i == 0:
            lg = eng.logits(0)           # logits of the last prompt token
            eng.sample(0)
        else:
            eng.decode([0])              # forwards the pending token and samples the next from its logits
            lg = eng.logits(0)
        top = np.partition(lg, -2)[-2:]
        margins.append(float(top[1] - top[0]))
    res.append((min(margins), ps, float(np.median(margins)), float(np.abs(lg).max())))
res.sort(reverse=True)
print(f"{arch} {ftype} model_seed {mseed} output_std {ostd}: {n_seeds} prompts in {time.time() - t0:.0f}s")
for m, ps, med, mx in res[:12]:
    print(f"prompt_seed {1000 + ps} min_margin {m:.4f} median {med:.3f} max|logit| {mx:.1f}")
