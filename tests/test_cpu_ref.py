"""The C restatement of ggml's CPU path (oracle/cpu_ref.c, the timed CPU baseline) against the
float32 oracle: equal up to the int8-activation noise ggml accepts (SURVEY.md 7 hard part #3)."""
import os

import numpy as np
import pytest


@pytest.mark.parametrize("arch,ftype,seed", [("tiny", "Q4_K_M", 1), ("tiny128", "Q4_0", 2), ("tiny", "Q8_0", 3), ("tiny128", "Q5_K", 4)])
def test_cpu_ref_close_to_float_oracle(tmp_models, arch, ftype, seed):
    from llama_cpp import synth
    from oracle.cpu_ref import CpuRef
    from oracle.llama_ref import RefModel
    path = os.path.join(tmp_models, f"cr_{arch}_{ftype}.gguf")
    synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=0.1)
    toks = np.random.default_rng(0).integers(0, 700, size=10)
    ref = RefModel(path).forward(toks)
    c = CpuRef(path, 64)
    for i, t in enumerate(toks):
        lg = c.eval(int(t))
        # int8 activations: a few percent of the logit spread; the arg-max agrees when the margin is clear
        assert np.abs(lg - ref[i]).max() < 0.05 * ref[i].std() * 4 + 0.05, i
    assert c.threads >= 1
    c.close()


@pytest.mark.parametrize("arch,ftype,seed", [("tiny", "Q4_K_M", 1), ("tiny128", "Q4_0", 2), ("tiny", "Q8_0", 3)])
def test_vnni_dots_equal_the_scalar_loops(tmp_models, arch, ftype, seed):
    """The AVX-512 VNNI block dots compute the same integers as the scalar loops: logits identical bit for bit."""
    from llama_cpp import synth
    from oracle.cpu_ref import CpuRef
    path = os.path.join(tmp_models, f"cr_{arch}_{ftype}.gguf")
    if not os.path.exists(path):
        synth.write_synthetic_model(path, arch, ftype, seed=seed, output_std=0.1)
    c = CpuRef(path, 32)
    if not c.has_vnni:
        c.close()
        pytest.skip("no AVX-512 VNNI on this CPU: only the scalar loops are built")
    toks = [int(t) for t in np.random.default_rng(1).integers(0, 700, size=6)]
    out = {}
    for mode in (True, False):
        c.set_simd(mode)
        c.reset()
        out[mode] = np.stack([c.eval(t) for t in toks])
    c.set_simd(True)
    c.close()
    assert np.array_equal(out[True].view(np.uint32), out[False].view(np.uint32))


def test_exact_mode_matches_the_float_oracle_tightly(tmp_models):
    """cr_set_exact: the same graph with float activations -- the checker bench.py uses on the 32-layer benchmark model,
    where the int8-activation noise of the timed mode is ~10% of the logit spread. Here: 1e-3 of the logit std."""
    from llama_cpp import synth
    from oracle.cpu_ref import CpuRef
    from oracle.llama_ref import RefModel
    path = os.path.join(tmp_models, "cr_exact.gguf")
    synth.write_synthetic_model(path, "tiny", "Q4_K_M", seed=1, output_std=0.1)
    toks = np.random.default_rng(0).integers(0, 700, size=8)
    ref = RefModel(path).forward(toks)
    c = CpuRef(path, 64)
    c.set_exact(True)
    try:
        for i, t in enumerate(toks):
            lg = c.eval(int(t))
            assert np.abs(lg - ref[i]).max() < 2e-3 * ref[i].std(), i
    finally:
        c.set_exact(False)
        c.close()
