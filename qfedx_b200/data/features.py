"""Batched, device-resident feature reduction for the quantum pipeline.

Torch restatements (batched, run on whatever device the data lives on) of the
reference's per-image NumPy helpers:

* ``downsample_images``  <- downsample_image  /root/reference/src/QFed/testEncoder.py:20-40
* ``pool_features``      <- pool_to_n_features /root/reference/src/QFed/qAngle.py:9-24
* ``angle_features``     <- the min-max * pi mapping in angle_encode /root/reference/src/QFed/qAngle.py:35-41
* ``amplitude_normalize``<- normalize_for_amplitude /root/reference/src/QFed/qAmplitude.py:11-22
* ``PCAReducer``         <- the PCA-to-k-features step that exists only as a plan
                            (/root/reference/ROADMAP.md:17-19)
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

__all__ = ["downsample_images", "pool_features", "angle_features", "amplitude_normalize",
           "PCAReducer", "images_to_angles", "images_to_amplitudes"]


def _edges(n_in: int, n_out: int):
    lo = [int(math.floor(i * n_in / n_out)) for i in range(n_out)]
    hi = [int(math.floor((i + 1) * n_in / n_out)) for i in range(n_out)]
    hi = [min(n_in, l + 1) if h <= l else h for l, h in zip(lo, hi)]
    return lo, hi


def _block_mean_matrix(n_in: int, n_out: int, device, dtype) -> torch.Tensor:
    """[n_out, n_in] averaging matrix with floor boundaries."""
    lo, hi = _edges(n_in, n_out)
    m = torch.zeros(n_out, n_in, device=device, dtype=dtype)
    for i, (l, h) in enumerate(zip(lo, hi)):
        m[i, l:h] = 1.0 / (h - l)
    return m


def downsample_images(imgs: torch.Tensor, out_shape=(4, 4)) -> torch.Tensor:
    """[..., H, W] -> [..., oh, ow] block means (28x28 -> 4x4 gives exact 7x7 means).
    Two small matmuls instead of a Python double loop per image."""
    h, w = imgs.shape[-2:]
    ah = _block_mean_matrix(h, out_shape[0], imgs.device, imgs.dtype)
    aw = _block_mean_matrix(w, out_shape[1], imgs.device, imgs.dtype)
    return ah @ imgs @ aw.transpose(0, 1)


def pool_features(vec: torch.Tensor, n_features: int) -> torch.Tensor:
    """[..., L] -> [..., n]: contiguous-chunk means, last chunk takes the tail; zero-pad if n >= L."""
    size = vec.shape[-1]
    if n_features >= size:
        out = vec.new_zeros(*vec.shape[:-1], n_features)
        out[..., :size] = vec
        return out
    chunk = size // n_features
    head = vec[..., : chunk * (n_features - 1)].reshape(*vec.shape[:-1], n_features - 1, chunk).mean(-1)
    tail = vec[..., chunk * (n_features - 1):].mean(-1, keepdim=True)
    return torch.cat([head, tail], dim=-1)


def angle_features(feats: torch.Tensor, n_qubits: int | None = None, scale: float = math.pi) -> torch.Tensor:
    """Per-sample min-max to [0,1] (constant vector -> zeros), times ``scale``."""
    if n_qubits is not None and feats.shape[-1] != n_qubits:
        feats = pool_features(feats, n_qubits)
    lo = feats.amin(-1, keepdim=True)
    span = feats.amax(-1, keepdim=True) - lo
    normed = torch.where(span > 0, (feats - lo) / torch.where(span > 0, span, torch.ones_like(span)),
                         torch.zeros_like(feats))
    return normed * scale


def amplitude_normalize(vec: torch.Tensor) -> torch.Tensor:
    """Unit-L2 rows; an all-zero row becomes the uniform state."""
    nrm = vec.norm(dim=-1, keepdim=True)
    uniform = torch.full_like(vec, 1.0 / math.sqrt(vec.shape[-1]))
    return torch.where(nrm > 0, vec / torch.where(nrm > 0, nrm, torch.ones_like(nrm)), uniform)


@dataclass
class PCAReducer:
    """PCA to k features fitted with an SVD on the device; ``state_dict`` round-trips so the
    fitted transformer can be saved next to the processed dataset."""
    k: int
    mean: torch.Tensor | None = None
    components: torch.Tensor | None = None  # [k, D]
    scale: torch.Tensor | None = None

    def fit(self, X: torch.Tensor, standardize: bool = True) -> "PCAReducer":
        X = X.reshape(X.shape[0], -1).to(torch.float32)
        self.mean = X.mean(0)
        Xc = X - self.mean
        self.scale = Xc.std(0).clamp_min(1e-6) if standardize else torch.ones_like(self.mean)
        Xc = Xc / self.scale
        # economy SVD of the covariance side that is smaller
        _, _, vh = torch.linalg.svd(Xc, full_matrices=False)
        comp = vh[: self.k]
        # deterministic sign: largest-magnitude coordinate of each component positive
        idx = comp.abs().argmax(dim=1)
        sign = torch.sign(comp[torch.arange(comp.shape[0]), idx])
        self.components = comp * sign[:, None]
        return self

    def transform(self, X: torch.Tensor) -> torch.Tensor:
        X = X.reshape(X.shape[0], -1).to(torch.float32)
        return ((X - self.mean) / self.scale) @ self.components.transpose(0, 1)

    def state_dict(self):
        return {"k": self.k, "mean": self.mean, "components": self.components, "scale": self.scale}

    @classmethod
    def from_state_dict(cls, sd):
        return cls(k=sd["k"], mean=sd["mean"], components=sd["components"], scale=sd["scale"])


def images_to_angles(X: torch.Tensor, n_qubits: int = 4, grid=(4, 4), scale: float = math.pi) -> torch.Tensor:
    """[N,1,28,28] (or [N,28,28]) -> [N, n_qubits] feature-map angles via the reference's
    demo path: block-mean to ``grid``, flatten, chunk-pool to n_qubits, min-max, * pi
    (/root/reference/src/QFed/testEncoder.py:94,124-125)."""
    if X.dim() == 4:
        X = X[:, 0]
    small = downsample_images(X, grid).reshape(X.shape[0], grid[0] * grid[1])   # explicit size: clients may be empty
    return angle_features(small, n_qubits, scale)


def images_to_amplitudes(X: torch.Tensor, n_qubits: int = 4) -> torch.Tensor:
    """[N,1,28,28] -> [N, 2^n] unit-norm real amplitudes (block-mean to a 2^n-pixel grid)."""
    if X.dim() == 4:
        X = X[:, 0]
    side_h = 1 << ((n_qubits + 1) // 2)
    side_w = 1 << (n_qubits // 2)
    small = downsample_images(X, (side_h, side_w)).reshape(X.shape[0], side_h * side_w)
    return amplitude_normalize(small)
