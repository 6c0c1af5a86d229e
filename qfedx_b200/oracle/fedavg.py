"""Float64 oracle for aggregation: FedAvg, DP clip + noise, pairwise masks.

* FedAvg by sample count over full state_dicts: /root/reference/src/CFed/Classical_FL.py:66-81
* clip to L2 norm C then add N(0, sigma^2 C^2 I): /root/reference/ROADMAP.md:50-51
* pairwise +m_ij / -m_ij masks that cancel in the sum: /root/reference/ROADMAP.md:52-55,137-138
"""
from __future__ import annotations

import numpy as np


def fedavg_state_dicts(updates):
    """updates: list of (dict name->ndarray, n_samples).  Weighted mean per key."""
    total = float(sum(n for _, n in updates))
    keys = list(updates[0][0].keys())
    out = {k: np.zeros_like(np.asarray(updates[0][0][k], dtype=np.float64)) for k in keys}
    for params, n in updates:
        w = n / total
        for k in keys:
            out[k] += w * np.asarray(params[k], dtype=np.float64)
    return out


def clip_l2(delta: np.ndarray, clip: float) -> np.ndarray:
    nrm = float(np.linalg.norm(delta))
    return delta * min(1.0, clip / max(nrm, 1e-30))


def aggregate_flat(deltas: np.ndarray, weights: np.ndarray, clip: float | None = None,
                   noise: np.ndarray | None = None, masks: np.ndarray | None = None,
                   participate: np.ndarray | None = None) -> np.ndarray:
    """deltas [K, D]; weights [K] (sample counts).  Returns sum_k w_k clip(delta_k) / sum w
    (+ noise, already scaled, + masks [K, D] which should cancel)."""
    deltas = np.asarray(deltas, dtype=np.float64)
    k = deltas.shape[0]
    part = np.ones(k, dtype=bool) if participate is None else np.asarray(participate, dtype=bool)
    w = np.asarray(weights, dtype=np.float64) * part
    w = w / w.sum()
    acc = np.zeros(deltas.shape[1])
    for i in range(k):
        if not part[i]:
            continue
        d = deltas[i] if clip is None else clip_l2(deltas[i], clip)
        acc += w[i] * d
        if masks is not None:
            acc += masks[i]
    if noise is not None:
        acc += noise
    return acc


def wrap_angle(x: np.ndarray) -> np.ndarray:
    """Wrap to [-pi, pi) (/root/reference/ROADMAP.md:37)."""
    return (np.asarray(x, dtype=np.float64) + np.pi) % (2 * np.pi) - np.pi
