"""SPSA gradient estimator (/root/reference/ROADMAP.md:38: "Adam + option for SPSA for high-cost
gradient reduction").  Two loss evaluations per step regardless of the parameter count:
g = (L(w + c D) - L(w - c D)) / (2 c) * D with D a Rademacher vector."""
from __future__ import annotations

import torch


@torch.no_grad()
def spsa_gradient(model, loss_closure, c: float = 0.1, generator=None):
    """Fills ``p.grad`` of every trainable parameter with the SPSA estimate; returns L(w)."""
    params = [p for p in model.parameters() if p.requires_grad]
    signs = [(torch.randint(0, 2, p.shape, device=p.device, generator=generator).float() * 2 - 1) for p in params]
    for p, s in zip(params, signs):
        p.add_(c * s)
    lp = loss_closure()
    for p, s in zip(params, signs):
        p.sub_(2 * c * s)
    lm = loss_closure()
    for p, s in zip(params, signs):
        p.add_(c * s)
    scale = (lp - lm) / (2 * c)
    for p, s in zip(params, signs):
        p.grad = scale * s
    return 0.5 * (lp + lm)
