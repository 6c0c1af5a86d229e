"""Rank <-> client mapping and the cross-GPU aggregation path.

``make_aggregator`` returns a callable with the signature of ``ops.aggregate.fedavg_aggregate``.
With one rank it is that function.  With several it is ``MultiGpuAggregator``: each rank reduces
its own clients and the partial sums are exchanged with in-kernel peer stores over NVLink inside
the same cooperative launch (csrc/fedavg_fused.cu, world > 1) - no NCCL call on the path.
``NcclAggregator`` is the measured baseline B1 (local torch reduction + ``all_reduce``).
"""
from __future__ import annotations

import ctypes
from typing import Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops.aggregate import MODES, AggResult, fedavg_aggregate
from ..fl import dp as _dp


def client_owner(client: int, n_clients: int, world: int) -> int:
    """Rank that hosts a client: contiguous blocks of ceil(K/world)."""
    per = (n_clients + world - 1) // world
    return client // per


def local_clients(rank: int, n_clients: int, world: int) -> list:
    per = (n_clients + world - 1) // world
    return list(range(rank * per, min(n_clients, (rank + 1) * per)))


class MultiGpuAggregator:
    """K7 across GPUs.  Small vectors use the one-barrier exchange (every rank pushes its partial to every peer);
    from ``two_shot_min_dim`` on a two-step exchange: through the NVSwitch reduction where the job has a multicast
    object (``exchange="nvls"``: the switch sums the W partials of a slice for its owner, who broadcasts the updated
    slice with one multicast store - D words per direction and rank), else reduce-scatter by peer stores + all-gather
    by peer loads (``"two_shot"``: 2 (W-1)/W D words).  ``exchange`` forces one of "one_shot", "two_shot", "nvls";
    ``two_shot`` is the older boolean spelling of the first two."""

    def __init__(self, rank: int, world: int, dim: int, device: torch.device, two_shot: bool | None = None,
                 two_shot_min_dim: int = 1 << 18, exchange: str | None = None):
        from .symm import MulticastBuffer, SymmetricBuffer
        self.rank, self.world, self.dim, self.device = rank, world, dim, device
        if exchange is None:
            if two_shot is None:
                exchange = "auto" if dim >= two_shot_min_dim else "one_shot"
            else:
                exchange = "two_shot" if two_shot else "one_shot"
        if exchange not in ("auto", "one_shot", "two_shot", "nvls"):
            raise ValueError(f"unknown exchange {exchange!r}")
        self.mc = None
        if exchange == "auto" and world < 4:
            # D (1 + 1/W) words per direction through the switch against 2 D (W-1)/W by peer stores and loads: at two
            # GPUs the peer exchange moves less (measured at 1e8 floats: 1.24 against 1.34 ms)
            exchange = "two_shot"
        if exchange in ("auto", "nvls"):
            # every rank takes the same branch: support is a property of the box, and the rendezvous is collective
            sup = torch.tensor([1 if MulticastBuffer.supported(device) else 0], device=device, dtype=torch.int32)
            dist.all_reduce(sup, op=dist.ReduceOp.MIN)
            if int(sup):
                self._out_off = (dim * 8 + 255) // 256 * 256
                mc = MulticastBuffer(self._out_off + dim * 4, device)
                if mc.available:
                    self.mc = mc
                    self.result_row = mc.local[self._out_off:self._out_off + dim * 4].view(torch.float32)
            if self.mc is None and exchange == "nvls":
                raise RuntimeError("exchange='nvls': no multicast (NVLS) support on this job's GPUs")
            exchange = "nvls" if self.mc is not None else "two_shot"
        self.exchange = exchange
        self.two_shot = exchange == "two_shot"
        self.flags = SymmetricBuffer(256, rank, world, device)
        self.flag_table = self.flags.peer_table()
        if exchange != "nvls":
            chunk = ((dim + world - 1) // world + 3) // 4 * 4      # words per rank slice (covers the 4-wide vector path)
            slot_bytes = ((world + 1) * chunk if self.two_shot else world * dim) * 8
            # two receive-slot sets (alternated by epoch parity)
            self.slots = [SymmetricBuffer(slot_bytes, rank, world, device) for _ in range(2)]
            self.slot_tables = [s.peer_table() for s in self.slots]
        self.epoch = 0
        self._scratch = torch.empty(4096, device=device, dtype=torch.float32)   # grown on demand (one float per local client)
        self._pair_keys = None

    def __call__(self, client_params, global_params, weights, participate=None, mode="delta", n_wrap=0, clip=0.0,
                 sigma=0.0, server_lr=1.0, seed=0, rnd=0, secagg=False, gids=None, out=None, pair_keys=None, **_):
        dev = self.device
        k_local = 0 if client_params is None else client_params.shape[0]
        k_global = weights.numel()
        if k_local > self._scratch.numel():      # the kernel clears and accumulates one squared norm per local client
            self._scratch = torch.empty(k_local, device=dev, dtype=torch.float32)
        w = weights.to(dev, torch.float32).contiguous()
        part = None if participate is None else torch.as_tensor(participate).to(dev, torch.int32).contiguous()
        g = global_params.contiguous().float()
        new = out if out is not None else torch.empty_like(g)
        gid_t = torch.as_tensor(np.asarray(gids if gids is not None else range(k_local)), dtype=torch.int32, device=dev)
        cp = client_params.contiguous().float() if k_local else g  # any valid pointer when there are no rows
        self.epoch += 1
        if secagg and pair_keys is not None and self._pair_keys is None:
            from ..ops.aggregate import pair_keys_tensor
            self._pair_keys = pair_keys_tensor(pair_keys, dev)
        pk = self._pair_keys if (secagg and pair_keys is not None) else None
        if self.exchange == "nvls":
            # one buffer, no parity: the launch's second cross-rank barrier separates a round's multicast reads from the
            # next round's writes.  out=self.result_row returns the broadcast row itself (no copy inside the launch).
            mc = self.mc
            _lib.call("qfx_fedavg_fused_multi_nvls", _lib.ptr(cp), _lib.ptr(g), _lib.ptr(new), _lib.ptr(w), _lib.ptr(part),
                      _lib.ptr(gid_t), _lib.c_int(k_local), _lib.c_int(k_global), _lib.c_int(self.dim),
                      _lib.c_int(MODES[mode]), _lib.c_int(n_wrap), _lib.c_float(clip or 0.0), _lib.c_float(sigma or 0.0),
                      _lib.c_float(server_lr), _lib.c_u64(seed & (2 ** 64 - 1)), _lib.c_int(rnd),
                      _lib.c_int(1 if secagg else 0), _lib.ptr(pk), _lib.ptr(None), _lib.ptr(self._scratch),
                      _lib.c_int(self.rank), _lib.c_int(self.world), ctypes.c_void_p(mc.local_ptr),
                      ctypes.c_void_p(mc.mc_ptr), ctypes.c_void_p(mc.local_ptr + self._out_off),
                      ctypes.c_void_p(mc.mc_ptr + self._out_off), _lib.ptr(self.flag_table), _lib.c_u32(self.epoch),
                      _lib.stream_ptr())
            return AggResult(new)
        tab = self.slot_tables[self.epoch & 1]
        _lib.call("qfx_fedavg_fused_multi_ex", _lib.ptr(cp), _lib.ptr(g), _lib.ptr(new), _lib.ptr(w), _lib.ptr(part),
                  _lib.ptr(gid_t), _lib.c_int(k_local), _lib.c_int(k_global), _lib.c_int(self.dim),
                  _lib.c_int(MODES[mode]), _lib.c_int(n_wrap), _lib.c_float(clip or 0.0), _lib.c_float(sigma or 0.0),
                  _lib.c_float(server_lr), _lib.c_u64(seed & (2 ** 64 - 1)), _lib.c_int(rnd),
                  _lib.c_int(1 if secagg else 0), _lib.ptr(pk), _lib.ptr(None), _lib.ptr(self._scratch),
                  _lib.c_int(self.rank), _lib.c_int(self.world), _lib.ptr(tab), _lib.ptr(self.flag_table),
                  _lib.c_u32(self.epoch), _lib.c_int(1 if self.two_shot else 0), _lib.stream_ptr())
        return AggResult(new)


class NcclAggregator:
    """Baseline B1: torch ops for clip/scale, NCCL all_reduce for the sum (not the product path)."""

    def __init__(self, rank, world, dim, device):
        self.rank, self.world, self.dim, self.device = rank, world, dim, device

    def __call__(self, client_params, global_params, weights, participate=None, mode="delta", n_wrap=0, clip=0.0,
                 sigma=0.0, server_lr=1.0, seed=0, rnd=0, secagg=False, gids=None, out=None, **_):
        dev = self.device
        w = weights.to(dev, torch.float32)
        part = torch.ones_like(w) if participate is None else torch.as_tensor(participate).to(dev, torch.float32)
        wn = w * part / (w * part).sum()
        acc = torch.zeros(self.dim, device=dev)
        if client_params is not None and client_params.shape[0]:
            rows = torch.as_tensor(np.asarray(gids), dtype=torch.long, device=dev)
            if mode == "weights":
                vals = client_params
            else:
                vals = _dp.wrap_angles(client_params - global_params[None], n_wrap)
                if clip and clip > 0:
                    vals = vals * _dp.clip_scale(vals, clip)[:, None]
                    if sigma and sigma > 0:
                        vals = vals + sigma * clip * torch.randn_like(vals)
            acc = (vals * (wn[rows] * part[rows])[:, None]).sum(0)
        if self.world > 1:
            dist.all_reduce(acc)
        new = global_params + server_lr * acc if mode == "delta" else acc
        return AggResult(new)


def make_aggregator(rank: int, world: int, gids: Sequence[int], all_sizes: torch.Tensor, dim: int, device,
                    kind: str = "fused"):
    if kind == "nccl":
        return NcclAggregator(rank, world, dim, device)
    if world == 1:
        return None  # the runner falls back to ops.aggregate.fedavg_aggregate
    return MultiGpuAggregator(rank, world, dim, device)
