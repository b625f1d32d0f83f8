"""One GEMM shape through b2_op_gemm repeatedly (for ncu): python scripts/gemm_micro.py TYPE N K T [reps]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llama-fastapi-k8s-gpu_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np
from llama_cpp import _cabi
from llama_cpp.synth import random_blocks
from util import TYPE_IDS, ptr
from oracle import ggml_blocks as gb
lib = _cabi.load_library()
tn, N, K, T = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
t = TYPE_IDS[tn]
w = random_blocks(tn, N * K // gb.BLOCK[t][0], 0.02, seed=1).reshape(-1)
x = np.random.default_rng(0).standard_normal((T, K)).astype(np.float32)
y = np.zeros((T, N), np.float32)
for _ in range(reps):
    rc = lib.b2_op_gemm(0, t, ptr(w), N, K, ptr(x), T, None, ptr(y), int(os.environ.get("SPLITS", "0")))
    assert rc == 0, lib.b2_last_error()
print("ok", float(np.abs(y).mean()))
