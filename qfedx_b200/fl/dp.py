"""Differential-privacy primitives on flat update vectors (torch, any device).

Clip each client's update to L2 norm C, add N(0, sigma^2 C^2 I) (/root/reference/ROADMAP.md:50-51).
These are the reference semantics of the fused CUDA aggregation (csrc/fedavg_fused.cu) and the
CPU path of ``qfedx_b200.ops.aggregate``.
"""
from __future__ import annotations

import math

import torch


def wrap_angles(delta: torch.Tensor, n_wrap: int | None = None) -> torch.Tensor:
    """Wrap the first n_wrap entries (all if None) to [-pi, pi) (/root/reference/ROADMAP.md:37)."""
    if n_wrap is None:
        n_wrap = delta.shape[-1]
    if n_wrap == 0:
        return delta
    head = delta[..., :n_wrap]
    head = head - 2 * math.pi * torch.floor((head + math.pi) / (2 * math.pi))
    return torch.cat([head, delta[..., n_wrap:]], dim=-1)


def clip_scale(deltas: torch.Tensor, clip: float) -> torch.Tensor:
    """[K, D] -> [K] factors min(1, C/||delta_k||)."""
    norms = deltas.norm(dim=1)
    return torch.where(norms > clip, clip / norms.clamp_min(1e-30), torch.ones_like(norms))


def clip_and_noise(deltas: torch.Tensor, clip: float, sigma: float, generator=None) -> torch.Tensor:
    out = deltas * clip_scale(deltas, clip)[:, None]
    if sigma > 0:
        out = out + sigma * clip * torch.randn(out.shape, device=out.device, dtype=out.dtype, generator=generator)
    return out
