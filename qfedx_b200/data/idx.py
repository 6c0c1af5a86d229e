"""MNIST idx readers/writers.

Same byte layout the reference consumes (/root/reference/src/CFed/Preprocess.py:11-20:
16-byte image header, 8-byte label header, uint8 payload), but the magic number and
the declared dimensions are validated instead of skipped, and a writer exists so
synthetic corpora can round-trip through the raw format.
"""
from __future__ import annotations

import struct
from pathlib import Path

import numpy as np

_MAGIC_IMAGES = 0x00000803
_MAGIC_LABELS = 0x00000801


def read_idx_images(path, check: bool = True) -> np.ndarray:
    raw = Path(path).read_bytes()
    if len(raw) < 16:
        raise ValueError(f"{path}: too short for an idx3 header")
    magic, count, rows, cols = struct.unpack(">IIII", raw[:16])
    if check and magic != _MAGIC_IMAGES:
        raise ValueError(f"{path}: bad idx3 magic 0x{magic:08x}")
    body = np.frombuffer(raw, dtype=np.uint8, offset=16)
    if check and body.size != count * rows * cols:
        raise ValueError(f"{path}: header says {count}x{rows}x{cols}, payload has {body.size} bytes")
    return body.reshape(-1, rows if check else 28, cols if check else 28)


def read_idx_labels(path, check: bool = True) -> np.ndarray:
    raw = Path(path).read_bytes()
    if len(raw) < 8:
        raise ValueError(f"{path}: too short for an idx1 header")
    magic, count = struct.unpack(">II", raw[:8])
    if check and magic != _MAGIC_LABELS:
        raise ValueError(f"{path}: bad idx1 magic 0x{magic:08x}")
    body = np.frombuffer(raw, dtype=np.uint8, offset=8)
    if check and body.size != count:
        raise ValueError(f"{path}: header says {count} labels, payload has {body.size}")
    return body


def write_idx_images(path, images: np.ndarray) -> None:
    images = np.ascontiguousarray(images, dtype=np.uint8)
    n, r, c = images.shape
    Path(path).write_bytes(struct.pack(">IIII", _MAGIC_IMAGES, n, r, c) + images.tobytes())


def write_idx_labels(path, labels: np.ndarray) -> None:
    labels = np.ascontiguousarray(labels, dtype=np.uint8)
    Path(path).write_bytes(struct.pack(">II", _MAGIC_LABELS, labels.size) + labels.tobytes())
