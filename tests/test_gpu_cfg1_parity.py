"""The designated CPU-vs-GPU parity run (BASELINE.json configs[0]): TinyLlama-1.1B-arch Q4_0
synthetic GGUF, 128-token prompt, 128 greedy tokens.

Protocol (hard part #2): the engine decodes greedily; the float32 oracle is then run teacher-forced
over the engine's own tokens in ONE batched pass, so a near-tie cannot hide the steps after it -- at every one of the
128 steps the oracle's arg-max is compared with the engine's token and the full logit vectors are compared. Random-init
logits are nearly flat, so the fixture (model seed, prompt seed, output_std) was screened on the GPU
(scripts/screen_margin_gpu.py, 400 prompts) for the largest minimum top-2 margin; the test re-measures that margin
with the oracle and records it.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ARCH, FTYPE, MODEL_SEED, OUTPUT_STD, PROMPT_SEED = "tinyllama-1.1b", "Q4_0", 1, 0.1, 1005
N_PROMPT = N_NEW = 128
SCREENED_MIN_MARGIN = 0.071      # engine-side figure from the screening run (gpurun_out/scr_01.log, round 2)


def test_cfg1_128_prompt_128_greedy_tokens_match_oracle(native_lib, tmp_models):
    from llama_cpp import synth
    from llama_cpp.engine import Engine, SamplingSettings
    from llama_cpp.gguf_file import GGUFFile
    from oracle.llama_ref import RefModel
    path = os.path.join(tmp_models, "cfg1.gguf")
    synth.write_synthetic_model(path, ARCH, FTYPE, seed=MODEL_SEED, output_std=OUTPUT_STD, with_vocab=False)
    eng = Engine(GGUFFile(path), n_ctx=N_PROMPT + N_NEW + 16, max_seqs=1)
    V = eng.dims.n_vocab
    prompt = [int(t) for t in np.random.default_rng(PROMPT_SEED).integers(0, V, N_PROMPT)]
    eng.ensure_pages(0, N_PROMPT + N_NEW + 1)
    eng.seq_begin(0, 0, [], SamplingSettings(temperature=0.0, repeat_penalty=1.0))
    eng.prefill(0, prompt)                         # tensor-core prompt pass
    got, logits = [], []
    for i in range(N_NEW):
        if i == 0:
            logits.append(eng.logits(0))
            got.append(eng.sample(0))
        else:
            got.append(eng.decode([0])[0])         # mat-vec decode step (CUDA-graph replay)
            logits.append(eng.logits(0))
    eng.close()
    ref = RefModel(path)
    want = ref.forward(prompt + got[:-1])[N_PROMPT - 1:]            # [128, V]: teacher-forced on the engine's tokens
    assert want.shape[0] == N_NEW
    top2 = np.partition(want, -2, axis=1)[:, -2:]
    margin = top2[:, 1] - top2[:, 0]
    oracle_tok = want.argmax(axis=1)
    err = np.array([np.abs(l - w).max() for l, w in zip(logits, want)])
    sigma = float(want.std())
    mism = [i for i in range(N_NEW) if oracle_tok[i] != got[i]]
    print(f"cfg1 parity: min oracle margin {margin.min():.4f} (step {int(margin.argmin())}), max |logit diff| {err.max():.4f} "
          f"(= {err.max() / sigma:.2e} of the logit std {sigma:.2f}), mismatching steps {mism}")
    # every step whose oracle margin is clear must agree; with this fixture that is every step
    for i in mism:
        assert margin[i] < 4 * err.max(), f"step {i}: engine {got[i]} != oracle {int(oracle_tok[i])} with margin {margin[i]:.4f}"
    assert margin.min() > 0.5 * SCREENED_MIN_MARGIN, "fixture drifted: re-run scripts/screen_margin_gpu.py"
    assert not mism, f"greedy tokens differ at steps {mism} (margins {[float(margin[i]) for i in mism]})"
    # Logits. The 128 prompt tokens go through the f16-operand tensor-core GEMM (operands rounded to 11 bits, the
    # precision class of the reference's own cuBLAS f16 prompt path), so every later step attends to K/V rows that carry
    # that rounding; the decode steps themselves run the int8x2 mat-vec. Measured on B200 (round 2): max |diff| 0.065 =
    # 1.3e-2 of the logit std 4.93, all 128 tokens identical, smallest oracle margin 0.077. Bar: 2e-2 of the logit std.
    assert err.max() <= 2e-2 * sigma, (err.max(), sigma)
