import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200"))
from llama_cpp import _cabi
from llama_cpp.synth import random_blocks
lib = _cabi.load_library(); P = ctypes.c_void_p
tname, tid, K, N = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
be = 32 if tid in (2, 8) else 256
w = random_blocks(tname, N * K // be, 0.02, 1).reshape(-1)
x = np.random.default_rng(0).standard_normal(K).astype(np.float32); y = np.zeros(N, np.float32)
for c in [int(v) for v in sys.argv[5].split(",")]:
    rc = lib.b2_op_matvec(0, tid, w.ctypes.data_as(P), N, K, x.ctypes.data_as(P), None, 0.0, None, y.ctypes.data_as(P), 2, c)
    print("ctas", c, "rc", rc, lib.b2_last_error(), float(np.abs(y).mean()), flush=True)
