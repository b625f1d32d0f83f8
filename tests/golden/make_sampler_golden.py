"""Writes tests/golden/sampler_draws.json: outputs of libstdc++'s std::mt19937 + std::discrete_distribution
(oracle/draw_ref.cpp, compiled here with g++) on seeded probability vectors. The fixture pins
oracle/sampler_ref.py (`draw`, `MT19937Uniforms`) to the very library routine llama.cpp's llama_sample_token
calls; it is committed so the comparison also runs where there is no compiler.

usage: python tests/golden/make_sampler_golden.py"""
import json
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def build_draw_ref() -> str:
    out_dir = os.path.join(ROOT, "oracle", "_ref")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "draw_ref")
    src = os.path.join(ROOT, "oracle", "draw_ref.cpp")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        subprocess.run(["g++", "-O2", "-o", exe, src], check=True)
    return exe


def make_cases():
    rng = np.random.default_rng(20240607)
    cases = []
    for i, n in enumerate([1, 2, 3, 5, 8, 17, 40, 40, 40, 40, 64, 200]):
        kind = i % 4
        if kind == 0:
            p = rng.random(n)
        elif kind == 1:                      # softmax of spread-out logits: a few dominant entries, like real candidate sets
            l = rng.standard_normal(n) * 3.0
            p = np.exp(l - l.max())
        elif kind == 2:                      # some exact zeros and ties
            p = np.round(rng.random(n) * 4) / 4
            p[0] = max(p[0], 0.25)
        else:
            p = np.exp(-np.arange(n, dtype=np.float64) * 0.35)
        p = (p / p.sum()).astype(np.float32)
        cases.append({"seed": int(rng.integers(0, 2 ** 32)), "n_draws": 24, "probs": [float(x) for x in p]})
    return cases


def run_draw_ref(exe, cases):
    text = [str(len(cases))]
    for c in cases:
        text.append(f"{c['seed']} {c['n_draws']} {len(c['probs'])} " + " ".join(np.format_float_scientific(np.float32(x), unique=True) for x in c["probs"]))
    r = subprocess.run([exe], input="\n".join(text) + "\n", capture_output=True, text=True, check=True)
    lines = r.stdout.strip().split("\n")
    assert len(lines) == len(cases)
    return [[int(v) for v in ln.split()] for ln in lines]


if __name__ == "__main__":
    cases = make_cases()
    draws = run_draw_ref(build_draw_ref(), cases)
    for c, d in zip(cases, draws):
        c["draws"] = d
    with open(os.path.join(HERE, "sampler_draws.json"), "w") as f:
        json.dump({"source": "oracle/draw_ref.cpp (g++ libstdc++ std::mt19937 + std::discrete_distribution<>)", "cases": cases}, f)
    print("wrote", len(cases), "cases")
