"""Micro-benchmark of the mat-vec kernel on the layer shapes of Llama-3-8B (run under
`ncu --metrics gpu__time_duration.sum` to read kernel times)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200"))
from llama_cpp import _cabi
from llama_cpp.synth import random_blocks
lib = _cabi.load_library()
P = ctypes.c_void_p
which = sys.argv[1] if len(sys.argv) > 1 else "gate_up"
ctas = [int(c) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
planes = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rng = np.random.default_rng(0)
def ptr(a): return a.ctypes.data_as(P)
if which == "gate_up":
    K, N = 4096, 14336
    wg = random_blocks("Q4_K", N * K // 256, 0.02, 1).reshape(-1); wu = random_blocks("Q4_K", N * K // 256, 0.02, 2).reshape(-1)
    x = rng.standard_normal(K).astype(np.float32); nw = np.ones(K, np.float32); y = np.zeros(N, np.float32)
    for c in ctas:
        for _ in range(2):
            rc = lib.b2_op_gate_up(0, 12, ptr(wg), 12, ptr(wu), N, K, ptr(x), ptr(nw), 1e-5, ptr(y), c)
        print("gate_up ctas", c, "rc", rc, float(np.abs(y).mean()))
else:
    tname, tid, K, N = {"down6": ("Q6_K", 14, 14336, 4096), "down4": ("Q4_K", 12, 14336, 4096), "head": ("Q6_K", 14, 4096, 128256),
                        "o": ("Q4_K", 12, 4096, 4096), "q8": ("Q8_0", 8, 4096, 14336)}[which]
    w = random_blocks(tname, N * K // 256 if tid > 8 else N * K // 32, 0.02, 1).reshape(-1)
    x = rng.standard_normal(K).astype(np.float32); y = np.zeros(N, np.float32)
    for c in ctas:
        for _ in range(2):
            rc = lib.b2_op_matvec(0, tid, ptr(w), N, K, ptr(x), None, 0.0, None, ptr(y), planes, c)
        print(which, "ctas", c, "rc", rc, float(np.abs(y).mean()))
