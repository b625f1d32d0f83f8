"""Device-timed streaming rate of the decode mat-vec kernel on the layer shapes of the benchmark models
(b2_bench_matvec: synthetic weights made on the GPU, launches back to back, CUDA events).
usage: python scripts/mv_micro.py [iters] [pdl 0|1] [shape ...]   -- prints us per launch and GB/s per shape."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200"))
from llama_cpp import _cabi
lib = _cabi.load_library()
BB = {2: (32, 18), 8: (32, 34), 12: (256, 144), 13: (256, 176), 14: (256, 210)}
# name: (mode, type, n_rows, K)
SHAPES = {
    "qkv_q4k": (0, 12, 6144, 4096), "o_q4k": (1, 12, 4096, 4096), "gateup_q4k": (3, 12, 14336, 4096),
    "down_q4k": (1, 12, 4096, 14336), "down_q6k": (1, 14, 4096, 14336), "head_q6k": (0, 14, 128256, 4096),
    "gateup_q8": (3, 8, 14336, 4096), "down_q8": (1, 8, 4096, 14336), "gateup_q40": (3, 2, 5632, 2048),
    "down70_q6k": (1, 14, 8192, 28672), "gateup70_q4k": (3, 12, 28672, 8192),
}
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pdl = int(sys.argv[2]) if len(sys.argv) > 2 else 1
names = sys.argv[3:] or list(SHAPES)
n_ctas = int(os.environ.get("MV_CTAS", "0"))
for planes in ((2, 1) if not os.environ.get("MV_PLANES") else (int(os.environ["MV_PLANES"]),)):
    for n in names:
        mode, t, N, K = SHAPES[n]
        be, bb = BB[t]
        nbytes = N * (K // be) * bb * (2 if mode == 3 else 1)
        copies = max(2, int(400e6 // nbytes) + 1)
        us = ctypes.c_float(0)
        rc = lib.b2_bench_matvec(0, mode, t, N, K, copies, iters, n_ctas, planes, pdl, ctypes.byref(us))
        if rc:
            print(n, "rc", rc, lib.b2_last_error().decode()); continue
        print(f"{n:14s} planes {planes} pdl {pdl} ctas {n_ctas} {os.environ.get('TAG', '')} {us.value:8.2f} us/launch  {nbytes / us.value / 1e3:8.1f} GB/s  ({nbytes / 1e6:.1f} MB, {copies} copies)", flush=True)
