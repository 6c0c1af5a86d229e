"""Iris as a 4-qubit federated task (/root/reference/ROADMAP.md:25: "Iris (4 features -> 4 qubits)").

The 150 x 4 table ships with scikit-learn (no download).  Features are min-max scaled per FEATURE over the training
split to feature-map angles in [0, pi] (the reference's angle encoder scales per sample, qAngle.py:35-41, which
would erase the between-sample differences a 4-number flower measurement carries), then partitioned over the
clients with the same IID / Dirichlet rules as the image data (data/partition.py)."""
from __future__ import annotations

import math
from typing import Sequence

import numpy as np
import torch

from .partition import partition
from .preprocess import Prepared, stratified_split


def load_iris(classes: Sequence[int] = (0, 1, 2)):
    """(X float32 [N, 4], y int64 [N] remapped to 0..C-1) for the requested species."""
    from sklearn.datasets import load_iris as _load
    d = _load()
    X, y = np.asarray(d.data, dtype=np.float32), np.asarray(d.target, dtype=np.int64)
    keep = np.isin(y, list(classes))
    X, y = X[keep], y[keep]
    remap = {c: i for i, c in enumerate(classes)}
    return torch.from_numpy(X), torch.from_numpy(np.vectorize(remap.get)(y).astype(np.int64))


def iris_angles(X: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
    return ((X - lo) / (hi - lo).clamp_min(1e-12)).clamp(0.0, 1.0) * math.pi


def prepare_iris(classes: Sequence[int] = (0, 1, 2), num_clients: int = 4, partition_type: str = "iid",
                 alpha: float = 0.5, test_split: float = 0.2, val_split: float = 0.1, seed: int = 42) -> Prepared:
    """Train / val / test splits and the client partition, features already as feature-map angles [N, 4]."""
    X, y = load_iris(classes)
    tr, te = stratified_split(y.numpy(), test_split, seed)
    Xte, yte = X[te], y[te]
    tr2, va = stratified_split(y[tr].numpy(), val_split, seed)
    Xtr, ytr, Xva, yva = X[tr][tr2], y[tr][tr2], X[tr][va], y[tr][va]
    lo, hi = Xtr.min(0).values, Xtr.max(0).values
    Atr, Ava, Ate = (iris_angles(t, lo, hi) for t in (Xtr, Xva, Xte))
    clients = partition(Atr, ytr, num_clients, partition_type, alpha)
    clients = [(a, b) if len(a) else (Atr[:1], ytr[:1]) for a, b in clients]
    return Prepared((Atr, ytr), (Ava, yva), (Ate, yte), clients)
