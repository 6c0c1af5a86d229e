"""Optional partition plots written to PNG files (no window, no display).

Counterpart of /root/reference/src/CFed/Preprocess.py:71-134.  matplotlib is not in the
build image, so the import is local and a missing matplotlib turns both calls into a
logged no-op; ``class_distribution_table`` gives the same information as text.
"""
from __future__ import annotations

import numpy as np

from .partition import partition_summary


def _mpl():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except ImportError:
        return None


def class_distribution_table(client_data) -> str:
    rows = partition_summary(client_data)
    classes = sorted({c for r in rows for c in r["classes"]})
    head = "client " + " ".join(f"{'d' + str(c):>7}" for c in classes) + "   total"
    body = [f"{r['client'] + 1:>6} " + " ".join(f"{r['classes'].get(c, 0):>7}" for c in classes) + f" {r['n']:>7}"
            for r in rows]
    return "\n".join([head] + body)


def visualize_client_data(client_data, save_path, samples_per_client=5) -> bool:
    plt = _mpl()
    if plt is None:
        print(f"matplotlib missing: skipped {save_path}")
        return False
    k = len(client_data)
    fig, axes = plt.subplots(k, samples_per_client, figsize=(2.4 * samples_per_client, 2.5 * k), squeeze=False)
    for i, (Xk, yk) in enumerate(client_data):
        for j in range(samples_per_client):
            ax = axes[i][j]
            ax.axis("off")
            if j < len(Xk):
                img = np.asarray(Xk[j]).reshape(28, 28)
                ax.imshow(img, cmap="gray")
                ax.set_title(f"label {int(yk[j])}")
    fig.suptitle("samples per client")
    fig.savefig(save_path, dpi=150, bbox_inches="tight")
    plt.close(fig)
    return True


def plot_class_distribution(client_data, save_path) -> bool:
    plt = _mpl()
    if plt is None:
        print(f"matplotlib missing: skipped {save_path}\n{class_distribution_table(client_data)}")
        return False
    rows = partition_summary(client_data)
    classes = sorted({c for r in rows for c in r["classes"]})
    names = [f"client {r['client'] + 1}" for r in rows]
    base = np.zeros(len(rows))
    fig, ax = plt.subplots(figsize=(10, 6))
    for c in classes:
        h = np.array([r["classes"].get(c, 0) for r in rows], dtype=float)
        ax.bar(names, h, bottom=base, label=f"digit {c}")
        base += h
    ax.set_ylabel("samples")
    ax.legend()
    fig.savefig(save_path, dpi=150, bbox_inches="tight")
    plt.close(fig)
    return True
