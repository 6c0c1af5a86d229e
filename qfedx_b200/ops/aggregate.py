"""Federated aggregation op: FedAvg + DP clip/noise + pairwise-mask secure aggregation + server
update.  CUDA tensors run the fused cooperative kernel K7 (csrc/fedavg_fused.cu); CPU tensors run
the restatement below (which is also the numerics reference in tests).

Semantics (see the kernel header for the formula): ``mode='weights'`` reproduces the reference's
federated_averaging (/root/reference/src/CFed/Classical_FL.py:66-81); ``mode='delta'`` averages
wrapped, clipped, noised deltas and applies ``theta_g += server_lr * update``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib

MODES = {"weights": 0, "delta": 1}


@dataclass
class AggResult:
    new_global: torch.Tensor
    norms: torch.Tensor | None = None
    masked: torch.Tensor | None = None


def _reference(client_params, global_params, weights, participate, mode, n_wrap, clip, sigma, server_lr,
               seed, rnd, secagg, want_masked, generator=None):
    from ..fl import dp as _dp          # local imports: qfedx_b200.fl imports this module
    from ..fl import secagg as _sa
    K, D = client_params.shape
    part = torch.ones(K, dtype=torch.bool) if participate is None else participate.bool().cpu()
    w = weights.double().cpu() * part.double()
    wn = (w / w.sum()).to(client_params.dtype).to(client_params.device)
    if mode == "weights":
        vals = client_params * wn[:, None]
        norms = None
    else:
        delta = _dp.wrap_angles(client_params - global_params[None], n_wrap)
        norms = delta.norm(dim=1)
        if clip and clip > 0:
            delta = delta * _dp.clip_scale(delta, clip)[:, None]
            if sigma and sigma > 0:
                delta = delta + sigma * clip * torch.randn(delta.shape, dtype=delta.dtype, device=delta.device,
                                                           generator=generator)
        vals = delta * wn[:, None]
    masked = None
    if secagg:
        idx = torch.nonzero(part).flatten().tolist()
        msgs = _sa.mask_messages(vals.detach().cpu().double().numpy(), seed, rnd, idx)
        upd = torch.from_numpy(_sa.unmask_sum(msgs)).to(client_params.dtype).to(client_params.device)
        if want_masked:
            masked = torch.from_numpy(msgs.view(np.int64).copy())
    else:
        upd = vals[part.to(vals.device)].sum(0)
    new = global_params + server_lr * upd if mode == "delta" else upd
    return AggResult(new, norms, masked)


def fedavg_aggregate(client_params: torch.Tensor, global_params: torch.Tensor, weights: torch.Tensor,
                     participate: torch.Tensor | None = None, mode: str = "delta", n_wrap: int = 0,
                     clip: float = 0.0, sigma: float = 0.0, server_lr: float = 1.0, seed: int = 0,
                     rnd: int = 0, secagg: bool = False, want_norms: bool = False, want_masked: bool = False,
                     out: torch.Tensor | None = None, scratch: torch.Tensor | None = None,
                     generator=None) -> AggResult:
    """client_params [K,D] fp32, global_params [D], weights [K] (sample counts)."""
    if mode not in MODES:
        raise ValueError(mode)
    if not _lib.use_cuda(client_params):
        res = _reference(client_params, global_params, weights, participate, mode, n_wrap, clip, sigma,
                         server_lr, seed, rnd, secagg, want_masked, generator)
        if out is not None:
            out.copy_(res.new_global)
            res.new_global = out
        return res
    K, D = client_params.shape
    dev = client_params.device
    cp = client_params.contiguous().float()
    g = global_params.contiguous().float()
    w = weights.to(dev, torch.float32).contiguous()
    part = None if participate is None else participate.to(dev, torch.int32).contiguous()
    new = out if out is not None else torch.empty_like(g)
    norms = torch.empty(K, device=dev, dtype=torch.float32) if (want_norms and mode == "delta") else None
    masked = torch.empty(K, D, device=dev, dtype=torch.int64) if (want_masked and secagg) else None
    if scratch is None:
        scratch = torch.empty(K, device=dev, dtype=torch.float32)
    _lib.call("qfx_fedavg_fused", _lib.ptr(cp), _lib.ptr(g), _lib.ptr(new), _lib.ptr(w), _lib.ptr(part),
              _lib.c_int(K), _lib.c_int(D), _lib.c_int(MODES[mode]), _lib.c_int(n_wrap), _lib.c_float(clip or 0.0),
              _lib.c_float(sigma or 0.0), _lib.c_float(server_lr), _lib.c_u64(seed & (2 ** 64 - 1)), _lib.c_int(rnd),
              _lib.c_int(1 if secagg else 0), _lib.ptr(norms), _lib.ptr(masked), _lib.ptr(scratch),
              _lib.stream_ptr())
    return AggResult(new, norms, masked)
