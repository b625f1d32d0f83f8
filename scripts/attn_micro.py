import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200"))
from llama_cpp import _cabi
lib = _cabi.load_library(); P = ctypes.c_void_p
T = int(sys.argv[1]) if len(sys.argv) > 1 else 600
splits = int(sys.argv[2]) if len(sys.argv) > 2 else 18
n_head, n_kv, hd, bt = 32, 8, 128, 64
rng = np.random.default_rng(0)
q = rng.standard_normal(n_head * hd).astype(np.float32)
k = (rng.standard_normal((T, n_kv, hd)) * 0.5).astype(np.float16); v = rng.standard_normal((T, n_kv, hd)).astype(np.float16)
nblk = (T + bt - 1) // bt
blocks = np.arange(nblk, dtype=np.int32); out = np.zeros(n_head * hd, np.float32)
for _ in range(3):
    rc = lib.b2_op_attention_decode(0, q.ctypes.data_as(P), k.ctypes.data_as(P), v.ctypes.data_as(P), T, n_head, n_kv, hd, bt,
                                    blocks.ctypes.data_as(P), nblk, splits, out.ctypes.data_as(P))
print("rc", rc, float(np.abs(out).mean()))
