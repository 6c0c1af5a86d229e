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


def pair_mask(seed: int, rnd: int, i: int, j: int, n_clients: int, dim: int, pair_keys=None) -> np.ndarray:
    """uint64 [dim] mask shared by clients i<j in round ``rnd``.  ``pair_keys`` [K, K] (uint64): the pair's secret from
    the key agreement keys the stream instead of the job seed (csrc/fedavg_fused.cu does the same)."""
    lo, hi = (i, j) if i < j else (j, i)
    base = _U(seed) if pair_keys is None else _U(pair_keys[lo, hi])
    k = key3(base ^ _U(TAG_MASK), rnd, lo * n_clients + hi)
    return stream_at(k, np.arange(dim, dtype=_U))


def uniform_bits(seed: int, tag: int, a: int, b: int, n: int) -> np.ndarray:
    """float64 [n] uniforms in [0,1) from the shared stream (client sampling, dropouts)."""
    k = key3(_U(seed) ^ _U(tag), a, b)
    return (stream_at(k, np.arange(n, dtype=_U)) >> _U(11)).astype(np.float64) / float(1 << 53)


def _feistel_round(x: int, k: int) -> int:
    x = (x ^ k) & 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x85EBCA6B) & 0xFFFFFFFF
    x ^= x >> 13
    x = (x * 0xC2B2AE35) & 0xFFFFFFFF
    return x ^ (x >> 16)


def feistel_permutation(n: int, key: int) -> np.ndarray:
    """Host mirror of feistel_perm (csrc/common.cuh): the keyed permutation of [0, n) that the training kernels
    use to visit a client's samples in a fresh order every epoch without storing a permutation."""
    if n <= 1:
        return np.zeros(n, dtype=np.int64)
    bits = 1
    while (1 << bits) < n:
        bits += 1
    a = bits >> 1
    b = bits - a
    key &= (1 << 64) - 1
    k0, k1 = key & 0xFFFFFFFF, key >> 32
    out = np.empty(n, dtype=np.int64)
    for j in range(n):
        x = j
        while True:
            l, r = x & ((1 << a) - 1), x >> a
            la, lb = a, b
            for rnd in range(4):
                k = ((k1 if rnd & 1 else k0) + 0x9E3779B9 * rnd) & 0xFFFFFFFF
                f = _feistel_round(r, k) & ((1 << la) - 1)
                l, r = r, l ^ f
                la, lb = lb, la
            x = (r << la) | l
            if x < n:
                break
        out[j] = x
    return out


def shuffle_key(seed: int, rnd: int) -> int:
    """Non-zero 64-bit key of the sample shuffle of one round."""
    k = int(key3(_U(seed) ^ _U(0x5348554646), rnd, 0))
    return k or 1
