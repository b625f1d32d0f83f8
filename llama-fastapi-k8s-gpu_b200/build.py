"""Build the in-tree native libraries.

  lib/libb200gguf.so       CUDA kernels + C ABI (include/b200_gguf.h), nvcc, sm_100a only
  lib/libb200_hosttest.so  host build of the block arithmetic (CPU self-test, g++)

nvcc cross-compiles without a GPU; the .so files are git-ignored but travel to the GPU box.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("B2_LIB_DIR") or os.path.join(HERE, "lib")   # B2_LIB_DIR: experiment builds (with B2_NVCC_EXTRA)
CUDA_SOURCES = ["engine.cu", "matvec.cu", "matvec_extra.cu", "attention.cu", "attention_mma.cu", "gemm.cu", "prefill.cu", "misc.cu", "sampler.cu", "tp.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _newer(target: str, sources) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIB, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b200_gguf.h")]
    out = os.path.join(LIB, "libb200gguf.so")
    extra = os.environ.get("B2_NVCC_EXTRA", "").split()
    if force or extra or _newer(out, deps):      # experiment flags: always rebuild (the up-to-date check knows nothing of flags)
        headers = [d for d in deps if not d.endswith(".cu")]

        def compile_one(src):
            obj = os.path.join(LIB, src.replace(".cu", ".o"))
            path = os.path.join(CSRC, src)
            if not force and not extra and not _newer(obj, [path, *headers]):
                return obj, 0, ""
            cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", path, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode == 0:
                with open(os.path.join(LIB, src + ".ptxas.log"), "w") as f:
                    f.write(r.stderr)
            return obj, r.returncode, r.stdout + r.stderr

        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(CUDA_SOURCES), os.cpu_count() or 4)) as ex:
            results = list(ex.map(compile_one, CUDA_SOURCES))
        objs = []
        for src, (obj, rc, log) in zip(CUDA_SOURCES, results):
            if verbose or rc:
                sys.stderr.write(log)
            if rc:
                raise RuntimeError(f"nvcc failed on {src}")
            objs.append(obj)
        cmd = [nvcc, "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    host = os.path.join(LIB, "libb200_hosttest.so")
    hsrc = os.path.join(CSRC, "host_selftest.cpp")
    if force or _newer(host, [hsrc, os.path.join(CSRC, "quant_blocks.cuh")]):
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", hsrc, "-o", host], check=True)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
