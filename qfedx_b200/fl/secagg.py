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


def client_mask(seed, rnd, k, participants, n_clients, dim) -> np.ndarray:
    """Net uint64 mask client k adds this round against the given participant set."""
    acc = np.zeros(dim, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in participants:
            if j == k:
                continue
            m = pair_mask(seed, rnd, k, j, n_clients, dim)
            acc = acc + m if k < j else acc - m
    return acc


def mask_messages(values: np.ndarray, seed: int, rnd: int, participants=None) -> np.ndarray:
    """values [K, D] float -> uint64 [K, D] masked fixed-point messages (rows of
    non-participants are zero)."""
    K, D = values.shape
    part = list(range(K)) if participants is None else [int(p) for p in participants]
    out = np.zeros((K, D), dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k in part:
            out[k] = encode_fixed(values[k]) + client_mask(seed, rnd, k, part, K, D)
    return out


def unmask_sum(messages: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        return decode_fixed(messages.sum(axis=0, dtype=np.uint64))


def unmask_dropped(total: np.ndarray, seed: int, rnd: int, survivors, dropped, n_clients: int) -> np.ndarray:
    """Remove from a uint64 sum over ``survivors`` the masks they added against ``dropped``
    clients (the server re-derives them from the revealed pair seeds)."""
    total = np.asarray(total, dtype=np.uint64).copy()
    dim = total.shape[0]
    with np.errstate(over="ignore"):
        for k in survivors:
            for j in dropped:
                m = pair_mask(seed, rnd, k, j, n_clients, dim)
                total = total - m if k < j else total + m
    return total
