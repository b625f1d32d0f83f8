"""ORACLE (test infrastructure, not product code) -- token sampling.

CPU restatement of the sampling chain llama-cpp-python 0.2.77 runs per token under
`create_chat_completion(temperature=1.2, top_p=0.9, frequency_penalty=0.7,
presence_penalty=0.8)` (/root/reference/api.py:55-63): `Llama.sample` ->
`_LlamaSamplingContext.sample` -> llama.cpp `llama_sample_repetition_penalties`,
`llama_sample_top_k`, `llama_sample_top_p`, `llama_sample_min_p`, `llama_sample_temp`,
`llama_sample_token` (un-vendored upstream). PARITY UNPINNED BY THE
REFERENCE: no llama_cpp here and no golden vectors in the reference repo; this restates the
published algorithm (tfs_z = typical_p = 1.0 are no-ops and are not modelled).

Independent artefacts that pin this module (tests/test_oracle_sampler.py):
  * `draw` + `MT19937Uniforms`: libstdc++'s own std::mt19937 + std::discrete_distribution<>, the two objects
    llama_sample_token uses, run by oracle/draw_ref.cpp -> tests/golden/sampler_draws.json (bit-exact indices).
    One difference is deliberate and documented: with a single candidate libstdc++ returns 0 WITHOUT drawing from
    the generator, while this restatement (and the device sampler) always spends one uniform per token.
  * penalties / top-k / top-p / min-p / temperature: hand-worked known-answer vectors, including the fixed
    settings of api.py:59-62.

Order: penalties over the last `penalty_last_n` tokens -> (temperature <= 0: arg-max, first
maximum) | top-k -> softmax -> top-p (smallest prefix with cumulative p >= top_p) -> min-p
(keep logit >= max_logit + log(min_p)) -> logits / temperature -> softmax ->
std::discrete_distribution draw with a uniform double from std::mt19937.
"""
from __future__ import annotations

import numpy as np


def apply_penalties(logits: np.ndarray, history, last_n: int, repeat: float, freq: float, present: float) -> np.ndarray:
    out = logits.astype(np.float32).copy()
    hist = list(history)[-min(last_n, 64):] if last_n > 0 else []
    if not hist or (repeat == 1.0 and freq == 0.0 and present == 0.0):
        return out
    counts = {}
    for t in hist:
        counts[t] = counts.get(t, 0) + 1
    for t, c in counts.items():
        if t < 0 or t >= len(out):
            continue
        l = out[t]
        l = np.float32(l * np.float32(repeat)) if l <= 0 else np.float32(l / np.float32(repeat))
        l = np.float32(l - (np.float32(c) * np.float32(freq) + np.float32(present)))
        out[t] = l
    return out


def candidates_after_filters(logits: np.ndarray, top_k: int, top_p: float, min_p: float, temperature: float):
    """Returns (ids, probs) of the final candidate set, sorted by logit descending (ties: low id first)."""
    V = len(logits)
    k = max(1, min(top_k, V))
    order = np.lexsort((np.arange(V), -logits.astype(np.float64)))[:k]
    l = logits[order].astype(np.float32)
    n = k
    if top_p < 1.0:
        p = np.exp((l - l[0]).astype(np.float32)).astype(np.float32)
        p = (p / np.float32(p.sum(dtype=np.float32))).astype(np.float32)
        cum = np.float32(0)
        last = n
        for i in range(n):
            cum = np.float32(cum + p[i])
            if cum >= np.float32(top_p):
                last = i + 1
                break
        n = last
    if min_p > 0.0 and n > 1:
        thr = np.float32(l[0] + np.float32(np.log(np.float32(min_p))))
        i = 1
        while i < n and not (l[i] < thr):
            i += 1
        n = i
    l = (l[:n] / np.float32(temperature)).astype(np.float32)
    e = np.exp((l - l[0]).astype(np.float32)).astype(np.float32)
    p = (e / np.float32(e.sum(dtype=np.float32))).astype(np.float32)
    return order[:n], p


def draw(probs: np.ndarray, u: float) -> int:
    """std::discrete_distribution: normalise in double, cumulative sums, last forced to 1,
    index = lower_bound(cp, u)."""
    p = probs.astype(np.float64)
    p = p / p.sum()
    cp = np.cumsum(p)
    cp[-1] = 1.0
    return int(np.searchsorted(cp, u, side="left"))


def sample(logits, history, u, temperature=0.8, top_k=40, top_p=0.95, min_p=0.05, repeat_penalty=1.1,
           frequency_penalty=0.0, presence_penalty=0.0, penalty_last_n=64) -> int:
    l = apply_penalties(np.asarray(logits, dtype=np.float32), history, penalty_last_n, repeat_penalty,
                        frequency_penalty, presence_penalty)
    if temperature <= 0.0 or top_k == 1:
        return int(np.argmax(l))
    ids, p = candidates_after_filters(l, top_k, top_p, min_p, temperature)
    return int(ids[draw(p, u)])


class MT19937Uniforms:
    """u = std::generate_canonical<double, 53>(std::mt19937(seed)): two 32-bit draws per double,
    (r0 + r1 * 2^32) / 2^64, as libstdc++ computes it for std::discrete_distribution."""

    def __init__(self, seed: int):
        bg = np.random.MT19937()
        bg._legacy_seeding(int(seed) & 0xFFFFFFFF)
        self._bg = bg

    def take(self, n: int) -> np.ndarray:
        raw = self._bg.random_raw(2 * n).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        r0 = raw[0::2].astype(np.float64)
        r1 = raw[1::2].astype(np.float64)
        u = (r0 + r1 * 4294967296.0) / 18446744073709551616.0
        return np.minimum(u, np.nextafter(1.0, 0.0))
