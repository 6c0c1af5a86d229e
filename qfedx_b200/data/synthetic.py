"""Synthetic MNIST-shape data (there is no network and the reference's image files are
stripped from the mount).

``synthetic_mnist`` produces class-conditional images: each class has a fixed random
low-frequency template plus per-sample noise, so models can actually learn (loss goes
down, accuracy beats chance) while shapes, dtypes and value ranges match what
/root/reference/src/CFed/Preprocess.py:180-196 produces: X float32 [N,1,28,28] in [0,1],
y int64 [N].
"""
from __future__ import annotations

import numpy as np
import torch

# MNIST class counts for digits 0..9 in the train / test label files (public dataset facts;
# the {0,1,2} rows are re-derived from the reference's label files in BASELINE.md section 2).
MNIST_TRAIN_COUNTS = (5923, 6742, 5958, 6131, 5842, 5421, 5918, 6265, 5851, 5949)
MNIST_TEST_COUNTS = (980, 1135, 1032, 1010, 982, 892, 958, 1028, 974, 1009)


def _templates(num_classes: int, gen: torch.Generator) -> torch.Tensor:
    coarse = torch.rand(num_classes, 1, 7, 7, generator=gen)
    t = torch.nn.functional.interpolate(coarse, size=(28, 28), mode="bilinear", align_corners=False)
    t = (t - t.amin(dim=(2, 3), keepdim=True)) / (t.amax(dim=(2, 3), keepdim=True) - t.amin(dim=(2, 3), keepdim=True))
    return t


def synthetic_mnist(n: int, num_classes: int = 3, seed: int = 0, noise: float = 0.25,
                    class_counts=None, contrast: float = 1.0, label_noise: float = 0.0):
    """Return (X float32 [n,1,28,28] in [0,1], y int64 [n]).

    ``contrast`` < 1 pulls the class templates towards their common mean and ``label_noise`` re-draws that fraction
    of the labels uniformly AFTER the images were made: together with a larger ``noise`` they give a task that a
    TinyCNN does not solve perfectly (accuracy ceiling about 1 - label_noise * (C-1)/C), which is what the
    convergence-parity tests need - the default task is separable within a round."""
    gen = torch.Generator().manual_seed(seed)
    if class_counts is not None:
        y = torch.cat([torch.full((int(c),), k, dtype=torch.long) for k, c in enumerate(class_counts)])
        y = y[torch.randperm(y.numel(), generator=gen)]
        n = y.numel()
    else:
        y = torch.randint(0, num_classes, (n,), generator=gen)
    tmpl = _templates(num_classes, torch.Generator().manual_seed(1234))
    if contrast != 1.0:
        mean = tmpl.mean(0, keepdim=True)
        tmpl = mean + contrast * (tmpl - mean)
    X = tmpl[y] + noise * torch.randn(n, 1, 28, 28, generator=gen)
    if label_noise > 0.0:
        flip = torch.rand(n, generator=gen) < label_noise
        y = torch.where(flip, torch.randint(0, num_classes, (n,), generator=gen), y)
    return X.clamp_(0.0, 1.0).contiguous(), y


def synthetic_uint8_corpus(digits=(0, 1, 2), seed: int = 0, scale: float = 1.0):
    """uint8 train/test image+label arrays with the real MNIST per-digit counts (optionally
    scaled down), suitable for writing idx files with qfedx_b200.data.idx."""
    out = {}
    for split, counts in (("train", MNIST_TRAIN_COUNTS), ("test", MNIST_TEST_COUNTS)):
        cc = [max(1, int(counts[d] * scale)) for d in digits]
        X, y = synthetic_mnist(0, len(digits), seed + (split == "test"), class_counts=cc)
        lab = np.asarray(digits, dtype=np.uint8)[y.numpy()]
        out[split] = ((X[:, 0].numpy() * 255.0).round().astype(np.uint8), lab)
    return out


def synthetic_images_for_labels(labels: np.ndarray, seed: int = 0, noise: float = 0.25) -> np.ndarray:
    """uint8 [N,28,28] class-conditional images for an existing label vector (used when only
    the label idx files are on disk, which is the state of the reference mount)."""
    gen = torch.Generator().manual_seed(seed)
    tmpl = _templates(10, torch.Generator().manual_seed(1234))
    y = torch.as_tensor(np.asarray(labels, dtype=np.int64))
    out = np.empty((y.numel(), 28, 28), dtype=np.uint8)
    for lo in range(0, y.numel(), 8192):
        yy = y[lo:lo + 8192]
        X = (tmpl[yy, 0] + noise * torch.randn(yy.numel(), 28, 28, generator=gen)).clamp_(0, 1)
        out[lo:lo + 8192] = (X * 255.0).round().to(torch.uint8).numpy()
    return out
