"""Client sampling and fault injection (/root/reference/ROADMAP.md:35,90-91,106).

``sample_clients`` picks the round's participants (a fixed-size fraction p, or Poisson sampling
with rate p - the assumption behind the accountant's subsampling amplification).
``inject_dropouts`` removes each selected client with probability ``drop_prob`` (fault injection);
the aggregation renormalises the weights over the survivors and secure-aggregation masks are
formed only among them (or repaired with ``secagg.unmask_dropped``).
Both are pure functions of (seed, round) so every rank derives the same sets with no exchange.
"""
from __future__ import annotations

import numpy as np

from .prng import uniform_bits

TAG_SAMPLE = 0x53414D50
TAG_DROP = 0x44524F50


def sample_clients(n_clients: int, fraction: float, rnd: int, seed: int = 0, poisson: bool = False,
                   allow_empty: bool = False) -> np.ndarray:
    """Sorted array of selected client ids.  Poisson sampling may come up empty: by default the client with the
    smallest draw is taken then (training always proceeds); ``allow_empty`` returns the empty set, which is what
    the sampled-Gaussian accountant assumes."""
    if fraction >= 1.0:
        return np.arange(n_clients)
    u = uniform_bits(seed, TAG_SAMPLE, rnd, 0, n_clients)
    if poisson:
        chosen = np.flatnonzero(u < fraction)
        return chosen if (chosen.size or allow_empty) else np.array([int(np.argmin(u))])
    m = max(1, int(round(fraction * n_clients)))
    return np.sort(np.argsort(u, kind="stable")[:m])


def inject_dropouts(selected: np.ndarray, drop_prob: float, rnd: int, seed: int = 0) -> np.ndarray:
    """Survivors among ``selected``; at least one client always survives."""
    if drop_prob <= 0:
        return selected
    u = uniform_bits(seed, TAG_DROP, rnd, 1, int(selected.max()) + 1)[selected]
    keep = selected[u >= drop_prob]
    return keep if keep.size else selected[:1]


def participation_mask(n_clients: int, participants) -> np.ndarray:
    m = np.zeros(n_clients, dtype=np.int32)
    m[np.asarray(participants, dtype=np.int64)] = 1
    return m
