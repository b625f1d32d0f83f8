"""Diagnostic: engine (graph / plain launches) vs the C restatement (VNNI / scalar) on the 8B benchmark model."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")]
import bench
from llama_cpp import _cabi
from llama_cpp.engine import Engine, SamplingSettings
from llama_cpp.gguf_file import GGUFFile
from oracle.cpu_ref import CpuRef
arch, ftype = (sys.argv[1], sys.argv[2]) if len(sys.argv) > 2 else ("llama-3-8b", "Q4_K_M")
path = bench.ensure_model(arch, ftype)
g = GGUFFile(path)
V = g.tensors["token_embd.weight"].shape[1]
prompt = [int(t) for t in np.random.default_rng(5).integers(0, V, size=6)]
outs = {}
for name, flags in (("graph", 0), ("plain", _cabi.B2_FLAG_NO_GRAPH | _cabi.B2_FLAG_NO_PDL)):
    eng = Engine(g, n_ctx=64, max_seqs=1, flags=flags)
    eng.ensure_pages(0, 32)
    eng.seq_begin(0, 0, [], SamplingSettings(temperature=0.0, repeat_penalty=1.0))
    eng.prefill(0, prompt)
    lg = [eng.logits(0)]; toks = [eng.sample(0)]
    for _ in range(3):
        toks.append(eng.decode([0])[0]); lg.append(eng.logits(0))
    outs[name] = (toks, lg)
    eng.close()
print("tokens graph", outs["graph"][0], "plain", outs["plain"][0])
print("graph vs plain max diff", [float(np.abs(a - b).max()) for a, b in zip(outs["graph"][1], outs["plain"][1])])
c = CpuRef(path, 64, n_threads=len(os.sched_getaffinity(0)))
res = {}
for mode in (True, "exact"):
    c.set_simd(True); c.set_exact(mode == "exact"); c.reset()
    for t in prompt[:-1]:
        c.eval(t, False)
    seq = [prompt[-1]] + outs["graph"][0][:-1]
    res[mode] = [c.eval(int(t)) for t in seq]
c.set_simd(True)
c.set_exact(False)
print("cpu int8 vs exact max diff", [float(np.abs(a - b).max()) for a, b in zip(res[True], res["exact"])])
for i, (gl, cl) in enumerate(zip(outs["graph"][1], res["exact"])):
    d = np.abs(gl - cl)
    j = int(d.argmax())
    print(f"step {i}: std {cl.std():.2f} max diff {d.max():.3f} at vocab {j} (gpu {gl[j]:.3f} cpu {cl[j]:.3f}); median {np.median(d):.4f} p99 {np.percentile(d, 99):.4f} "
          f"n>1.0: {int((d > 1).sum())}; argmax gpu {int(gl.argmax())} cpu {int(cl.argmax())}; corr {np.corrcoef(gl, cl)[0, 1]:.5f}")
    big = np.where(d > 1)[0]
    if len(big):
        print("   rows with big diff: first", big[:12], "... mod 2:", np.bincount(big % 2, minlength=2), "range", big.min(), big.max())
