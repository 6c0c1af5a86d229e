"""Dataset preparation: raw idx -> filtered, normalised, split, saved, partitioned.

Capability parity with /root/reference/src/CFed/Preprocess.py:137-228 and the on-disk
formats of SURVEY.md section 2.2:

* raw: ``{train,t10k}-{images.idx3,labels.idx1}-ubyte``
* processed: ``torch.save((X float32 [N,1,28,28] in [0,1], y int64 [N]), "{train,val,test}.pt")``

Differences on purpose: client partitions are returned as torch tensors with int64 labels
(the reference returns NumPy uint8 and its training loop then fails);
labels can be remapped to 0..C-1 (``remap_labels``) so digit sets other than (0,1,2) work;
missing raw files fall back to a synthetic corpus when ``synthetic_fallback`` is set; plots
are optional file output.  ``load_processed`` reads the .pt files back, which the reference
never does.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import idx as _idx
from .partition import partition, partition_summary
from .synthetic import synthetic_images_for_labels, synthetic_uint8_corpus

RAW_FILES = {
    "train_images": "train-images.idx3-ubyte",
    "train_labels": "train-labels.idx1-ubyte",
    "test_images": "t10k-images.idx3-ubyte",
    "test_labels": "t10k-labels.idx1-ubyte",
}

TensorPair = Tuple[torch.Tensor, torch.Tensor]


@dataclass
class Prepared:
    train: TensorPair
    val: TensorPair
    test: TensorPair
    clients: List[TensorPair]

    def __iter__(self):  # unpack like the reference's 4-tuple
        return iter((self.train, self.val, self.test, self.clients))


def stratified_split(y: np.ndarray, test_size: float, seed: int = 42):
    """(train_idx, val_idx).  Uses sklearn's splitter when present so the split is the one the
    reference gets (``train_test_split(stratify=y, random_state=42)``); otherwise a
    per-class permutation split with the same proportions."""
    try:
        from sklearn.model_selection import train_test_split
        tr, va = train_test_split(np.arange(len(y)), test_size=test_size, stratify=y, random_state=seed)
        return np.asarray(tr), np.asarray(va)
    except ImportError:  # pragma: no cover - sklearn is in the image
        rng = np.random.RandomState(seed)
        tr, va = [], []
        for c in np.unique(y):
            m = rng.permutation(np.flatnonzero(y == c))
            k = int(round(len(m) * test_size))
            va.extend(m[:k]); tr.extend(m[k:])
        return rng.permutation(np.asarray(tr)), rng.permutation(np.asarray(va))


def filter_and_normalize(images_u8: np.ndarray, labels_u8: np.ndarray, digits, remap_labels=False):
    keep = np.isin(labels_u8, digits)
    X = (images_u8[keep].astype(np.float32) / 255.0)[:, None, :, :]
    y = labels_u8[keep]
    if remap_labels:
        lut = np.full(256, -1, dtype=np.int64)
        for new, d in enumerate(digits):
            lut[d] = new
        y = lut[y]
    return X, y


def load_raw(raw_folder, synthetic_fallback=False, digits=(0, 1, 2), synthetic_scale=1.0):
    """dict with train/test (images uint8 [N,28,28], labels uint8 [N]); None if files are
    missing and no fallback was asked for (the reference's behaviour)."""
    paths = {k: os.path.join(raw_folder, v) for k, v in RAW_FILES.items()}
    have_labels = all(os.path.exists(paths[k]) for k in ("train_labels", "test_labels"))
    have_images = all(os.path.exists(paths[k]) for k in ("train_images", "test_images"))
    if have_labels and have_images:
        return {"train": (_idx.read_idx_images(paths["train_images"]), _idx.read_idx_labels(paths["train_labels"])),
                "test": (_idx.read_idx_images(paths["test_images"]), _idx.read_idx_labels(paths["test_labels"]))}
    if not synthetic_fallback:
        return None
    if have_labels and synthetic_scale == 1.0:
        # keep the real label files (they ship with the reference) and draw matching images
        out = {}
        for split, key in (("train", "train_labels"), ("test", "test_labels")):
            lab = _idx.read_idx_labels(paths[key])
            out[split] = (synthetic_images_for_labels(lab, seed=7 + (split == "test")), lab)
        return out
    return synthetic_uint8_corpus(digits, scale=synthetic_scale)


def preprocess_mnist(raw_folder="./dataset/raw", processed_folder="./dataset/processed",
                     digits=(0, 1, 2), val_split=0.1, num_clients=4, partition_type="iid",
                     alpha=0.5, remap_labels=False, synthetic_fallback=False, synthetic_scale=1.0,
                     save=True, plots=False, results_folder="./results", verbose=True) -> Optional[Prepared]:
    log = print if verbose else (lambda *a, **k: None)
    raw = load_raw(raw_folder, synthetic_fallback, digits, synthetic_scale)
    if raw is None:
        log(f"raw idx files not found under {raw_folder}")
        return None
    Xtr, ytr = filter_and_normalize(*raw["train"], digits, remap_labels)
    Xte, yte = filter_and_normalize(*raw["test"], digits, remap_labels)
    tr_idx, va_idx = stratified_split(ytr, val_split, seed=42)
    to_pair = lambda X, y: (torch.from_numpy(np.ascontiguousarray(X)), torch.as_tensor(np.asarray(y), dtype=torch.long))
    train, val, test = to_pair(Xtr[tr_idx], ytr[tr_idx]), to_pair(Xtr[va_idx], ytr[va_idx]), to_pair(Xte, yte)
    if save:
        Path(processed_folder).mkdir(parents=True, exist_ok=True)
        for name, pair in (("train", train), ("val", val), ("test", test)):
            torch.save(pair, os.path.join(processed_folder, f"{name}.pt"))
    for name, pair in (("train", train), ("val", val), ("test", test)):
        cls, cnt = np.unique(pair[1].numpy(), return_counts=True)
        log(f"  {name}: {tuple(pair[0].shape)}  " + ", ".join(f"digit {c}: {m}" for c, m in zip(cls, cnt)))
    clients = partition(train[0], train[1], num_clients, partition_type, alpha)
    for row in partition_summary(clients):
        log(f"  client {row['client'] + 1}: {row['n']} samples {row['classes']}")
    if plots:
        from .plots import plot_class_distribution, visualize_client_data
        Path(results_folder).mkdir(parents=True, exist_ok=True)
        visualize_client_data(clients, os.path.join(results_folder, "client_samples.png"))
        plot_class_distribution(clients, os.path.join(results_folder, "class_distribution.png"))
    return Prepared(train, val, test, clients)


def load_processed(processed_folder="./dataset/processed", splits=("train", "val", "test")):
    """Read the ``{split}.pt`` tuples written by ``preprocess_mnist`` (or by the reference)."""
    return {s: tuple(torch.load(os.path.join(processed_folder, f"{s}.pt"), weights_only=True)) for s in splits}
