"""Counter-based PRNG shared bit-for-bit with the CUDA kernels (csrc/common.cuh: mix64, key3,
stream_at).  Both parties of a secure-aggregation pair derive the same mask stream from
(seed, round, pair) with no exchange, and the server can re-derive the masks of a client that
dropped (/root/reference/ROADMAP.md:52-55: "PRG-based masks ... seed exchange API")."""
from __future__ import annotations

import numpy as np

_U = np.uint64
TAG_NOISE = 0x4E4F495345
TAG_MASK = 0x4D41534B53


def mix64(z):
    z = np.asarray(z, dtype=_U)
    with np.errstate(over="ignore"):
        z = z + _U(0x9E3779B97F4A7C15)
        z = (z ^ (z >> _U(30))) * _U(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> _U(27))) * _U(0x94D049BB133111EB)
        return z ^ (z >> _U(31))


def key3(seed, a, b):
    with np.errstate(over="ignore"):
        return mix64(mix64(mix64(_U(seed)) ^ _U(a)) ^ (_U(b) * _U(0xD1342543DE82EF95)))


def stream_at(key, idx):
    with np.errstate(over="ignore"):
        return mix64(_U(key) ^ (np.asarray(idx, dtype=_U) * _U(0x9E3779B97F4A7C15)))


def pair_mask(seed: int, rnd: int, i: int, j: int, n_clients: int, dim: int) -> np.ndarray:
    """uint64 [dim] mask shared by clients i<j in round ``rnd``."""
    lo, hi = (i, j) if i < j else (j, i)
    k = key3(_U(seed) ^ _U(TAG_MASK), rnd, lo * n_clients + hi)
    return stream_at(k, np.arange(dim, dtype=_U))


def uniform_bits(seed: int, tag: int, a: int, b: int, n: int) -> np.ndarray:
    """float64 [n] uniforms in [0,1) from the shared stream (client sampling, dropouts)."""
    k = key3(_U(seed) ^ _U(tag), a, b)
    return (stream_at(k, np.arange(n, dtype=_U)) >> _U(11)).astype(np.float64) / float(1 << 53)
