"""Evaluation metrics beyond accuracy: ROC AUC (/root/reference/ROADMAP.md:111-113 lists "accuracy / AUC").

``roc_auc`` is the rank statistic (probability that a random positive scores above a random negative, ties count
half); ``auc_from_logits`` is the binary AUC for one-logit models and the macro one-vs-rest AUC otherwise."""
from __future__ import annotations

import torch


def roc_auc(scores: torch.Tensor, positive: torch.Tensor) -> float:
    s = scores.detach().double().flatten().cpu()
    pos = positive.detach().bool().flatten().cpu()
    n_pos, n_neg = int(pos.sum()), int((~pos).sum())
    if n_pos == 0 or n_neg == 0:
        return float("nan")
    order = torch.argsort(s)
    ranks = torch.empty_like(s)
    ranks[order] = torch.arange(1, s.numel() + 1, dtype=torch.double)
    # average the ranks of tied scores
    uniq, inv, cnt = torch.unique(s, return_inverse=True, return_counts=True)
    if uniq.numel() < s.numel():
        sums = torch.zeros(uniq.numel(), dtype=torch.double).index_add_(0, inv, ranks)
        ranks = (sums / cnt.double())[inv]
    return float((ranks[pos].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


def auc_from_logits(logits: torch.Tensor, y: torch.Tensor) -> float:
    if logits.dim() == 1 or logits.shape[1] == 1:
        return roc_auc(logits.reshape(-1), y == 1)
    vals = [roc_auc(logits[:, c], y == c) for c in range(logits.shape[1])]
    vals = [v for v in vals if v == v]
    return float(sum(vals) / len(vals)) if vals else float("nan")
