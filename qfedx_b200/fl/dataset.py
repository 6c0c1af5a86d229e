"""Device-resident federated dataset: every client's samples packed on one tensor with segment
offsets.  The reference re-uploads each client's whole dataset every round
(/root/reference/src/CFed/Classical_FL.py:47-48, 13 MB per client per round); here the data is
moved once and stays resident for the run."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import numpy as np
import torch


@dataclass
class FederatedDataset:
    X: torch.Tensor            # [N, ...] features (images, angles or amplitudes)
    y: torch.Tensor            # [N] labels
    offsets: torch.Tensor      # [K+1] int32 on the same device
    sizes: np.ndarray          # [K] host copy of the segment lengths

    @classmethod
    def from_clients(cls, clients: Sequence[Tuple[torch.Tensor, torch.Tensor]], device=None,
                     pin: bool = False) -> "FederatedDataset":
        sizes = np.array([len(c[0]) for c in clients], dtype=np.int64)
        X = torch.cat([torch.as_tensor(c[0]) for c in clients], 0)
        y = torch.cat([torch.as_tensor(c[1]).long() for c in clients], 0)
        if pin and torch.cuda.is_available():
            X, y = X.pin_memory(), y.pin_memory()
        if device is not None:
            X, y = X.to(device, non_blocking=True), y.to(device, non_blocking=True)
        off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int32, device=X.device)
        return cls(X, y, off, sizes)

    @property
    def num_clients(self) -> int:
        return len(self.sizes)

    @property
    def device(self):
        return self.X.device

    def client(self, k: int):
        lo, hi = int(self.sizes[:k].sum()), int(self.sizes[:k + 1].sum())
        return self.X[lo:hi], self.y[lo:hi]

    def map_features(self, fn) -> "FederatedDataset":
        return FederatedDataset(fn(self.X), self.y, self.offsets, self.sizes)

    def epoch_permutations(self, epochs: int, generator=None) -> torch.Tensor:
        """int32 [epochs, N]: row e holds, segment by segment, a random permutation of each
        client's global sample indices (one sort for all clients and epochs)."""
        N = self.X.shape[0]
        owner = torch.repeat_interleave(torch.arange(self.num_clients, device=self.device),
                                        torch.as_tensor(self.sizes, device=self.device))
        keys = torch.rand(epochs, N, device=self.device, generator=generator) + 2.0 * owner[None, :].float()
        return keys.argsort(dim=1).to(torch.int32).contiguous()
