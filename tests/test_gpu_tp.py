"""Tensor parallelism on real GPUs (needs two): the sharded group against a single-GPU engine on the same GGUF --
logits within the mat-vec path's bar, greedy tokens identical, the same tokens on every rank. The toy model covers the
mechanics, the one-layer 70B-shaped model the real row / column splits (K = 28672 / 2, 64 heads / 2)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("arch,ftype,world", [("tiny", "Q4_K_M", 2), ("wide70-1l", "Q4_K_M", 2)])
def test_tensor_parallel_group_matches_single_gpu(native_lib, arch, ftype, world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29600 + (os.getpid() % 300)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "tp_worker.py"), arch, ftype]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "tp-parity-ok" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
