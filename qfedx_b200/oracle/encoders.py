"""NumPy float64 restatements of the reference's encoder helpers (the oracle for
qfedx_b200.data.features and the CUDA feature-map prologue).

Behaviour pinned to:
* normalize_for_amplitude  /root/reference/src/QFed/qAmplitude.py:11-22
* amplitude_encode         /root/reference/src/QFed/qAmplitude.py:25-41 (state only; no gate synthesis)
* pool_to_n_features       /root/reference/src/QFed/qAngle.py:9-24
* angle_encode             /root/reference/src/QFed/qAngle.py:27-51 (angles + resulting product state)
* downsample_image         /root/reference/src/QFed/testEncoder.py:20-40
"""
from __future__ import annotations

import math

import numpy as np

from .statevector import product_state_from_angles


def normalize_for_amplitude(vec) -> np.ndarray:
    v = np.asarray(vec, dtype=np.float64).copy()
    nrm = math.sqrt(float(np.sum(v * v)))
    if nrm == 0.0:
        return np.full(v.shape, 1.0 / math.sqrt(v.size))
    return v / nrm


def amplitude_state(vec) -> np.ndarray:
    """complex128 statevector sum_i x_i |i> (what Statevector.from_instruction returns)."""
    v = np.asarray(vec, dtype=np.float64).reshape(-1)
    size = v.size
    if size == 0 or size & (size - 1):
        raise ValueError(f"amplitude encoding needs a power-of-two length, got {size}")
    return normalize_for_amplitude(v).astype(np.complex128)


def pool_to_n_features(vec, n_features: int) -> np.ndarray:
    v = np.asarray(vec, dtype=np.float64).reshape(-1)
    size = v.size
    out = np.zeros(n_features)
    if n_features >= size:
        out[:size] = v
        return out
    chunk = size // n_features
    for i in range(n_features):
        lo = i * chunk
        hi = size if i == n_features - 1 else lo + chunk
        out[i] = v[lo:hi].mean()
    return out


def angle_features(features, n_qubits: int | None = None) -> np.ndarray:
    """Rotation angles in [0, pi]: pool to n_qubits, per-sample min-max, times pi."""
    f = np.asarray(features, dtype=np.float64).reshape(-1)
    if n_qubits is None:
        n_qubits = f.size
    if f.size != n_qubits:
        f = pool_to_n_features(f, n_qubits)
    span = f.max() - f.min()
    normed = np.zeros_like(f) if span == 0 else (f - f.min()) / span
    return normed * np.pi


def angle_state(features, n_qubits: int | None = None, basis: str = "ry") -> np.ndarray:
    return product_state_from_angles(angle_features(features, n_qubits), basis)


def downsample_image(img, out_shape=(4, 4)) -> np.ndarray:
    a = np.asarray(img, dtype=np.float64)
    in_h, in_w = a.shape
    out_h, out_w = out_shape
    out = np.zeros(out_shape)
    for i in range(out_h):
        h0 = int(math.floor(i * in_h / out_h))
        h1 = int(math.floor((i + 1) * in_h / out_h))
        if h1 <= h0:
            h1 = min(in_h, h0 + 1)
        for j in range(out_w):
            w0 = int(math.floor(j * in_w / out_w))
            w1 = int(math.floor((j + 1) * in_w / out_w))
            if w1 <= w0:
                w1 = min(in_w, w0 + 1)
            patch = a[h0:h1, w0:w1]
            out[i, j] = patch.mean() if patch.size else 0.0
    return out
