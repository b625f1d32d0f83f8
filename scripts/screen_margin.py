"""Search for a (seed, output_std) whose greedy continuation has a comfortable top-2 margin at every step, for the
designated parity run (BASELINE.json configs[0]: TinyLlama-1.1B-arch Q4_0, 128-token prompt + 128 greedy tokens).
Screening uses the fast C restatement (oracle/cpu_ref.c, int8 activations); the winner is re-checked with the float32
oracle by the test itself. usage: python scripts/screen_margin.py arch ftype output_std seed0 n_seeds [n_prompt n_new]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")]
from llama_cpp import synth
from oracle.cpu_ref import CpuRef

arch, ftype, ostd, seed0, n_seeds = sys.argv[1], sys.argv[2], float(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
n_prompt = int(sys.argv[6]) if len(sys.argv) > 6 else 128
n_new = int(sys.argv[7]) if len(sys.argv) > 7 else 128
V = synth.ARCHS[arch].n_vocab
for seed in range(seed0, seed0 + n_seeds):
    t0 = time.time()
    path = f"/tmp/screen_{arch}_{ftype}.gguf"
    synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=ostd, with_vocab=False)
    ref = CpuRef(path, n_prompt + n_new + 8)
    prompt = [int(t) for t in np.random.default_rng(seed).integers(0, V, n_prompt)]
    for t in prompt[:-1]:
        ref.eval(t, False)
    lg = ref.eval(prompt[-1], True)
    margins, toks = [], []
    for i in range(n_new):
        top = np.partition(lg, -2)[-2:]
        margins.append(float(top[1] - top[0]))
        tok = int(np.argmax(lg)); toks.append(tok)
        if i + 1 < n_new:
            lg = ref.eval(tok, True)
    ref.close()
    print(f"seed {seed} output_std {ostd} min_margin {min(margins):.4f} median {np.median(margins):.3f} distinct {len(set(toks))} "
          f"max|logit| {float(np.abs(lg).max()):.1f} ({time.time() - t0:.0f}s)", flush=True)
