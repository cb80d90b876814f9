"""Synthetic inputs of BASELINE.json's configs: bounded-Zipf id streams with a
fixed rank->row bijection so hot rows are scattered over the table.  numpy only."""
import math

import numpy as np

CONFIGS = {
    # name: (height V, width D, ids per step N, zipf a, optimizer)
    "C1_word2vec": (2073, 32, 4 * 100, 1.1, "sgd"),
    "C2_ctr_sgd": (1_000_000, 64, 4096 * 26, 1.1, "sgd"),
    "C3_adam": (10_000_000, 128, 16384 * 40, 1.1, "adam"),
    "C4_sharded_sgd": (100_000_000, 64, 4096 * 26, 1.1, "sgd"),
    "C5_dup_heavy": (50_000_000, 256, 65536 * 100, 1.3, "sgd"),
}

_MULT = 2654435761


def _coprime_multiplier(height):
    m = _MULT % height
    if m == 0:
        m = 1
    while math.gcd(m, height) != 1:
        m += 1
    return m


class ZipfIds:
    """P(rank r) ~ r^-a over ranks 1..V by inverse CDF; row = (rank*mult + 12345) mod V."""

    def __init__(self, height, a, seed=0):
        self.height = int(height)
        self.a = float(a)
        self.rng = np.random.Generator(np.random.PCG64(seed))
        w = np.arange(1, self.height + 1, dtype=np.float64) ** (-self.a)
        self.cdf = np.cumsum(w)
        self.cdf /= self.cdf[-1]
        self.mult = _coprime_multiplier(self.height)

    def sample(self, n):
        u = self.rng.random(int(n))
        rank = np.searchsorted(self.cdf, u, side="left").astype(np.int64) + 1
        rank = np.minimum(rank, self.height)
        return (rank * self.mult + 12345) % self.height


def table_values(height, d, seed=1234, dtype=np.float32, lo=-0.05, hi=0.05):
    rng = np.random.Generator(np.random.PCG64(seed))
    return (rng.random((height, d), dtype=np.float32) * (hi - lo) + lo).astype(dtype)


def grad_values(n, d, seed=4321, dtype=np.float32, scale=1e-2):
    rng = np.random.Generator(np.random.PCG64(seed))
    return (rng.standard_normal((n, d), dtype=np.float32) * scale).astype(dtype)


def algorithmic_bytes(n, d, u, optimizer, elem=4):
    """SURVEY.md §8d / BASELINE.md §3."""
    fwd = n * (8 + 2 * elem * d)
    if optimizer == "adam":
        upd = n * (8 + elem * d) + u * (6 * elem * d)
    else:
        upd = n * (8 + elem * d) + u * (2 * elem * d)
    return {"fwd": fwd, "upd": upd, "step": fwd + upd}
