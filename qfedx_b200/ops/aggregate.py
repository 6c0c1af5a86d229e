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
               seed, rnd, secagg, want_masked, generator=None, pair_keys=None, late_dropped=None):
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
        pk = None if pair_keys is None else np.asarray(pair_keys.cpu() if torch.is_tensor(pair_keys) else pair_keys).astype(np.uint64)
        late = [int(j) for j in (late_dropped or [])]
        # the protocol as it happens: survivors masked against everybody selected at send time; the messages of the
        # late drop-outs never arrive, the survivors reveal those pair secrets and the server strips the orphaned masks
        msgs = _sa.mask_messages(vals.detach().cpu().double().numpy(), seed, rnd, idx, pk, mask_against=idx + late)
        with np.errstate(over="ignore"):
            total = msgs.sum(axis=0, dtype=np.uint64)
        if late:
            total = _sa.unmask_dropped(total, seed, rnd, idx, late, K, pk)
        upd = torch.from_numpy(_sa.decode_fixed(total)).to(client_params.dtype).to(client_params.device)
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
                     generator=None, pair_keys=None, late_dropped=None) -> AggResult:
    """client_params [K,D] fp32, global_params [D], weights [K] (sample counts).

    ``pair_keys`` [K,K] uint64 (fl.secagg.KeyAgreement.table()): per-pair mask secrets instead of the job seed.
    ``late_dropped``: clients (not in ``participate``) that dropped after the others had masked against them; the CPU
    path plays the unmasking protocol, the CUDA kernel forms masks among the survivors only - the same sum."""
    if mode not in MODES:
        raise ValueError(mode)
    if not _lib.use_cuda(client_params):
        res = _reference(client_params, global_params, weights, participate, mode, n_wrap, clip, sigma,
                         server_lr, seed, rnd, secagg, want_masked, generator, pair_keys, late_dropped)
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
    if secagg and pair_keys is not None:
        pk = pair_keys_tensor(pair_keys, dev)
        _lib.call("qfx_fedavg_fused_pairkeys", _lib.ptr(cp), _lib.ptr(g), _lib.ptr(new), _lib.ptr(w), _lib.ptr(part),
                  _lib.c_int(K), _lib.c_int(D), _lib.c_int(MODES[mode]), _lib.c_int(n_wrap), _lib.c_float(clip or 0.0),
                  _lib.c_float(sigma or 0.0), _lib.c_float(server_lr), _lib.c_u64(seed & (2 ** 64 - 1)), _lib.c_int(rnd),
                  _lib.ptr(pk), _lib.ptr(norms), _lib.ptr(masked), _lib.ptr(scratch), _lib.stream_ptr())
        return AggResult(new, norms, masked)
    _lib.call("qfx_fedavg_fused", _lib.ptr(cp), _lib.ptr(g), _lib.ptr(new), _lib.ptr(w), _lib.ptr(part),
              _lib.c_int(K), _lib.c_int(D), _lib.c_int(MODES[mode]), _lib.c_int(n_wrap), _lib.c_float(clip or 0.0),
              _lib.c_float(sigma or 0.0), _lib.c_float(server_lr), _lib.c_u64(seed & (2 ** 64 - 1)), _lib.c_int(rnd),
              _lib.c_int(1 if secagg else 0), _lib.ptr(norms), _lib.ptr(masked), _lib.ptr(scratch),
              _lib.stream_ptr())
    return AggResult(new, norms, masked)


def pair_keys_tensor(pair_keys, device) -> torch.Tensor:
    """uint64 pair-secret table -> contiguous int64 device tensor (bit pattern preserved) for the kernels."""
    if torch.is_tensor(pair_keys):
        return pair_keys.to(device=device, dtype=torch.int64).contiguous()
    arr = np.ascontiguousarray(np.asarray(pair_keys, dtype=np.uint64)).view(np.int64)
    return torch.from_numpy(arr.copy()).to(device)
