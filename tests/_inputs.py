"""Seeded synthetic inputs for the tests (numpy only; the libstdc++-exact generators of
generate_inputs.cpp live in the reference shim, see _libs.ref_generate)."""
import numpy as np


def zipf(n, alphabet=1 << 20, alpha=1.0, seed=0, offset=1):
    """Zipf(alpha) ranks in [offset, offset + alphabet) by inverse CDF."""
    rng = np.random.default_rng(seed)
    w = 1.0 / np.arange(1, alphabet + 1, dtype=np.float64) ** alpha
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    r = np.searchsorted(cdf, rng.random(n), side="left")
    return (np.minimum(r, alphabet - 1) + offset).astype(np.uint32)


def geometric(n, p=0.01, seed=0):
    rng = np.random.default_rng(seed)
    return (rng.geometric(p, n) - 1).astype(np.uint32)


def uniform_bits(n, bits, seed=0):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 1 << bits, n, dtype=np.uint64).astype(np.uint32)


def wide(n, seed=0):
    """log-uniform magnitudes up to 2^30 - 1: exercises 0..3 raw bytes and every fold boundary."""
    rng = np.random.default_rng(seed)
    bits = rng.integers(0, 31, n)
    v = rng.integers(0, 1 << 30, n, dtype=np.uint64) >> (30 - bits).astype(np.uint64)
    edge = np.array([0, 1, 255, 256, 257, 511, 512, 65535, 65536, 65537, (1 << 24) - 1, 1 << 24, (1 << 24) + 1,
                     (1 << 30) - 1], dtype=np.uint64)
    v[: min(n, edge.size)] = edge[: min(n, edge.size)]
    return v.astype(np.uint32)


def make(kind, n, seed=0):
    if kind == "zipf20":
        return zipf(n, 1 << 20, 1.0, seed)
    if kind == "zipf16":
        return zipf(n, 1 << 16, 1.0, seed)
    if kind == "zipf12s":  # skewed, small alphabet
        return zipf(n, 1 << 12, 1.5, seed, offset=0)
    if kind == "geom":
        return geometric(n, 0.01, seed)
    if kind == "geom.2":
        return geometric(n, 0.2, seed)
    if kind == "u8":
        return uniform_bits(n, 8, seed)
    if kind == "u16":
        return uniform_bits(n, 16, seed)
    if kind == "wide":
        return wide(n, seed)
    raise KeyError(kind)
