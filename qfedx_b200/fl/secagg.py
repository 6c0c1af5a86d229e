"""Pairwise-mask secure aggregation in fixed point (/root/reference/ROADMAP.md:52-55,137-138).

Client i adds +m_ij for every peer j>i and -m_ij for every peer j<i to its fixed-point update;
all arithmetic is mod 2^64 so the masks cancel EXACTLY in the sum (float masks would not, and
would interact with DP noise - SURVEY.md 7.3).  ``mask_messages`` is the host restatement of
what csrc/fedavg_fused.cu does per element; ``unmask_dropped`` removes the residue left by
clients that dropped after their peers had already masked against them.
"""
from __future__ import annotations

import numpy as np

from .prng import pair_mask

FIXED_SCALE = float(1 << 32)


def encode_fixed(x: np.ndarray) -> np.ndarray:
    return np.rint(np.asarray(x, dtype=np.float64) * FIXED_SCALE).astype(np.int64).view(np.uint64)


def decode_fixed(q: np.ndarray) -> np.ndarray:
    return np.asarray(q, dtype=np.uint64).view(np.int64).astype(np.float64) / FIXED_SCALE


class KeyAgreement:
    """Registration-time seed exchange, simulated (/root/reference/ROADMAP.md:52-55: "seed exchange API (simulate DH key
    exchange during registration)").

    Every client draws a private exponent from its OWN random stream, publishes g^a mod p, the server relays the
    public keys, and clients i and j both arrive at g^(a_i a_j) mod p without it ever crossing the wire; its hash is the
    64-bit secret that keys the pair's mask stream.  A survivor can ``reveal`` the secret it shares with a client that
    dropped after masking, so that the server can strip the orphaned masks (``unmask_dropped``).
    p = 2^127 - 1 and the SplitMix-style hash are simulation-grade (a toy group and a non-cryptographic PRG): the
    protocol flow is real, the security level is not."""

    P = (1 << 127) - 1
    G = 3

    def __init__(self, n_clients: int, seed: int = 0):
        from .prng import key3, mix64
        self.n = int(n_clients)
        # private exponents: in a deployment each client draws its own; here stream k stands in for client k's RNG
        self._priv = [(int(key3(np.uint64(seed & (2 ** 64 - 1)) ^ np.uint64(0x4B455953), k, 0)) << 63 |
                       int(key3(np.uint64(seed & (2 ** 64 - 1)) ^ np.uint64(0x4B455953), k, 1))) % (self.P - 2) + 1
                      for k in range(self.n)]
        self.public = [pow(self.G, a, self.P) for a in self._priv]          # what registration sends to the server
        self._mix = mix64

    def shared_secret(self, me: int, peer: int) -> int:
        """What client ``me`` computes from its private exponent and ``peer``'s public key."""
        s = pow(self.public[peer], self._priv[me], self.P)
        return int(self._mix(np.uint64(s & (2 ** 64 - 1)) ^ self._mix(np.uint64(s >> 64))))

    def table(self) -> np.ndarray:
        """uint64 [K, K] symmetric table of pair secrets (what the simulation hands to the aggregation kernel)."""
        t = np.zeros((self.n, self.n), dtype=np.uint64)
        for i in range(self.n):
            for j in range(i + 1, self.n):
                t[i, j] = t[j, i] = np.uint64(self.shared_secret(i, j))
        return t

    def reveal(self, survivor: int, dropped: int) -> int:
        """The secret ``survivor`` discloses to the server about its pair with ``dropped``."""
        return self.shared_secret(survivor, dropped)


def client_mask(seed, rnd, k, participants, n_clients, dim, pair_keys=None) -> np.ndarray:
    """Net uint64 mask client k adds this round against the given participant set."""
    acc = np.zeros(dim, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in participants:
            if j == k:
                continue
            m = pair_mask(seed, rnd, k, j, n_clients, dim, pair_keys)
            acc = acc + m if k < j else acc - m
    return acc


def mask_messages(values: np.ndarray, seed: int, rnd: int, participants=None, pair_keys=None,
                  mask_against=None) -> np.ndarray:
    """values [K, D] float -> uint64 [K, D] masked fixed-point messages (rows of non-participants are zero).
    ``mask_against``: the set every participant masked against when it sent its message (defaults to the
    participants; larger when some of those clients dropped afterwards)."""
    K, D = values.shape
    part = list(range(K)) if participants is None else [int(p) for p in participants]
    against = part if mask_against is None else [int(p) for p in mask_against]
    out = np.zeros((K, D), dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k in part:
            out[k] = encode_fixed(values[k]) + client_mask(seed, rnd, k, against, K, D, pair_keys)
    return out


def unmask_sum(messages: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        return decode_fixed(messages.sum(axis=0, dtype=np.uint64))


def unmask_dropped(total: np.ndarray, seed: int, rnd: int, survivors, dropped, n_clients: int, pair_keys=None) -> np.ndarray:
    """Remove from a uint64 sum over ``survivors`` the masks they added against ``dropped``
    clients (the server re-derives them from the revealed pair seeds)."""
    total = np.asarray(total, dtype=np.uint64).copy()
    dim = total.shape[0]
    with np.errstate(over="ignore"):
        for k in survivors:
            for j in dropped:
                m = pair_mask(seed, rnd, k, j, n_clients, dim, pair_keys)
                total = total - m if k < j else total + m
    return total
