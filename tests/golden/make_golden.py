"""Generate the committed golden fixtures (run in the build container, where gguf-py and
transformers are importable; neither needs the GPU):

  blocks.npz      random quant blocks of every supported type + their expansion by
                  `gguf.quants.dequantize` (ggml's own numpy reference)  -> pins oracle/ggml_blocks.py
                  and the CUDA dequantise kernel (bit-exact)
  tiny_model.npz  logits of an INDEPENDENT implementation (transformers' GGUF loader ->
                  float32 LlamaForCausalLM, HF rotate-half RoPE on un-permuted weights) on the
                  seeded synthetic `tiny` and `tiny128` models -> pins oracle/llama_ref.py
                  (graph, GQA, RoPE pairing convention); the file's sha256 guards the generator.

The reference repo itself has no golden vectors.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

MODELS = [("tiny", "Q4_K_M", 1), ("tiny128", "Q4_0", 2), ("tiny", "Q8_0", 3), ("tiny128", "Q5_K", 4)]
TOKENS = np.random.default_rng(42).integers(0, 700, size=24)


def sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def main():
    from gguf.constants import GGMLQuantizationType as T
    from gguf.quants import dequantize
    from util import TYPE_IDS, finite_random_blocks
    rng = np.random.default_rng(2024)
    out = {}
    for tname in ("Q4_0", "Q8_0", "Q4_K", "Q5_K", "Q6_K"):
        raw = finite_random_blocks(tname, 64, rng)
        out[f"{tname}_raw"] = raw
        out[f"{tname}_f32"] = dequantize(raw, getattr(T, tname)).reshape(-1).astype(np.float32)
    rng2 = np.random.default_rng(2025)          # types added in round 2: their own stream, the entries above stay as they were
    for tname in ("Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K"):
        raw = finite_random_blocks(tname, 64, rng2)
        out[f"{tname}_raw"] = raw
        out[f"{tname}_f32"] = dequantize(raw, getattr(T, tname)).reshape(-1).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "blocks.npz"), **out)

    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    from transformers.modeling_gguf_pytorch_utils import load_gguf_checkpoint
    from llama_cpp import synth
    res = {"tokens": TOKENS}
    for arch, ftype, seed in MODELS:
        path = f"/tmp/golden_{arch}_{ftype}.gguf"
        synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=0.1)
        ck = load_gguf_checkpoint(path, return_tensors=False)
        cfg = LlamaConfig(**{k: v for k, v in ck["config"].items() if k != "model_type"})
        cfg.tie_word_embeddings = False
        with torch.device("meta"):
            dummy = LlamaForCausalLM(cfg)
        ck = load_gguf_checkpoint(path, return_tensors=True, model_to_load=dummy)
        hf = LlamaForCausalLM(cfg).float().eval()
        hf.load_state_dict(ck["tensors"], strict=True)
        with torch.no_grad():
            logits = hf(torch.tensor(TOKENS)[None]).logits[0].numpy().astype(np.float32)
        key = f"{arch}_{ftype}_{seed}"
        res[key + "_logits"] = logits
        res[key + "_sha256"] = np.frombuffer(bytes.fromhex(sha256(path)), dtype=np.uint8)
        print(key, logits.shape, float(logits.std()))
    np.savez_compressed(os.path.join(HERE, "tiny_model.npz"), **res)


if __name__ == "__main__":
    main()
