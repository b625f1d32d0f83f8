"""Host side of sampling: the uniform random stream for the device sampler.

llama.cpp draws the token with `std::discrete_distribution` over a `std::mt19937` seeded from
the context seed (reached from /root/reference/api.py:55-63 via llama-cpp-python `Llama.sample`;
SURVEY.md 2.2 H8). The device sampler consumes one double per token; this produces those doubles
the way libstdc++ does: `generate_canonical<double, 53>` = two 32-bit MT19937 outputs,
(r0 + r1 * 2^32) / 2^64.
"""
from __future__ import annotations

import os

import numpy as np


class UniformStream:
    def __init__(self, seed: int | None = None):
        if seed is None or seed < 0 or seed == 0xFFFFFFFF:      # LLAMA_DEFAULT_SEED: pick one at random
            seed = int.from_bytes(os.urandom(4), "little")
        self.seed = int(seed) & 0xFFFFFFFF
        self._bg = np.random.MT19937()
        self._bg._legacy_seeding(self.seed)                     # init_genrand(seed) == std::mt19937(seed)

    def take(self, n: int) -> np.ndarray:
        raw = self._bg.random_raw(2 * n).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        u = (raw[0::2].astype(np.float64) + raw[1::2].astype(np.float64) * 4294967296.0) / 18446744073709551616.0
        return np.minimum(u, np.nextafter(1.0, 0.0))
