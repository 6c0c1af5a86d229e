"""Client partitioners: IID and Dirichlet non-IID.

Index-level restatement of /root/reference/src/CFed/Preprocess.py:23-68 with the same
remainder rules, so that with the same seeds the split is the same one:

* IID: one shuffle, ``len // K`` per client, the last client takes the tail.
* Dirichlet: per class, shuffle that class's indices, draw p ~ Dir(alpha * 1_K), give
  client k ``int(len * p_k)`` samples, the last client takes what is left.  A client's
  indices are ordered class by class.

Unlike the reference (which iterates ``range(num_classes)`` and silently yields empty
clients for label sets such as (3, 5, 8)), the classes are
taken from the labels that actually occur.  The partitioners return index arrays; use
``take`` to slice tensors or arrays, which keeps torch tensors as torch tensors (the
reference returns NumPy and then crashes in ``client_update``).
"""
from __future__ import annotations

import random as _random
from typing import List, Sequence

import numpy as np

__all__ = ["iid_indices", "dirichlet_indices", "take", "partition", "partition_summary"]


def _py_rng(rng):
    return _random if rng is None else rng


def iid_indices(n_items: int, num_clients: int, rng=None) -> List[np.ndarray]:
    order = list(range(n_items))
    _py_rng(rng).shuffle(order)
    per = n_items // num_clients
    bounds = [k * per for k in range(num_clients)] + [n_items]
    return [np.asarray(order[bounds[k]:bounds[k + 1]], dtype=np.int64) for k in range(num_clients)]


def dirichlet_indices(labels: Sequence[int], num_clients: int, alpha: float = 0.5,
                      rng=None, np_rng=None) -> List[np.ndarray]:
    labels = np.asarray(labels)
    draw = np.random.dirichlet if np_rng is None else np_rng.dirichlet
    buckets: List[list] = [[] for _ in range(num_clients)]
    for cls in np.unique(labels):
        members = np.flatnonzero(labels == cls).tolist()
        _py_rng(rng).shuffle(members)
        share = draw([alpha] * num_clients)
        lo = 0
        for k in range(num_clients):
            hi = len(members) if k == num_clients - 1 else lo + int(len(members) * share[k])
            buckets[k].extend(members[lo:hi])
            lo = hi
    return [np.asarray(b, dtype=np.int64) for b in buckets]


def take(X, y, idx):
    """Slice (X, y) by an index array; works for numpy arrays and torch tensors."""
    try:
        import torch
        if isinstance(X, torch.Tensor):
            t = torch.as_tensor(idx, dtype=torch.long, device=X.device)
            return X.index_select(0, t), y.index_select(0, t)
    except ImportError:  # pragma: no cover
        pass
    return X[idx], y[idx]


def partition(X, y, num_clients: int, kind: str = "iid", alpha: float = 0.5, rng=None, np_rng=None):
    """Return ``[(X_k, y_k)]`` for ``kind`` in {'iid', 'non_iid'}."""
    if kind == "iid":
        parts = iid_indices(len(X), num_clients, rng)
    elif kind in ("non_iid", "dirichlet"):
        lab = y.cpu().numpy() if hasattr(y, "cpu") else np.asarray(y)
        parts = dirichlet_indices(lab, num_clients, alpha, rng, np_rng)
    else:
        raise ValueError(f"unknown partition type {kind!r}")
    return [take(X, y, p) for p in parts]


def partition_summary(client_data) -> List[dict]:
    """Per-client sample count and class histogram (the numbers the reference prints at
    /root/reference/src/CFed/Preprocess.py:217-221)."""
    rows = []
    for k, (_, yk) in enumerate(client_data):
        lab = yk.cpu().numpy() if hasattr(yk, "cpu") else np.asarray(yk)
        cls, cnt = np.unique(lab, return_counts=True)
        rows.append({"client": k, "n": int(lab.size),
                     "classes": {int(c): int(m) for c, m in zip(cls, cnt)}})
    return rows
