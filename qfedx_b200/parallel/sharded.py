"""Sharded statevector (K6): an n-qubit state split over 2^g GPUs on the top g "global" qubits.

``plan_sharded`` is the host-side scheduler: it walks the gate list with a logical->physical qubit
map and emits alternating steps

  ('local', passes)  - gates whose qubits are all local, compiled into K5 passes on the shard
  ('swap',  ops)     - exchange the g global qubits with the top g local ones (an all-to-all of
                       contiguous slices) FUSED with the following gates on the swapped-in qubits

using the same dependency-aware scan as the pass compiler (gates on disjoint qubits commute).  It is a
pure function of (gates, n, g) and is tested on the CPU for every world size.

``ShardedState`` executes a plan.  ``loopback=True`` keeps all shards in one device's memory and runs
the ranks' kernels one after another (peer pointers aliased locally, no in-kernel barrier) so the
kernels and the plan can be checked on a single GPU; the multi-process mode maps peers' shards
through the symmetric heap and the kernel synchronises with system-scope flags - no NCCL on the path.
``nccl_swap_baseline`` is B1: the same exchange with ``torch.distributed.all_to_all_single``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

import numpy as np
import torch

from ..ops import _lib
from ..ops.large import TWO_QUBIT, Pass, compile_passes, expvals_z, run_passes

Gate = Tuple[int, int, int, int]


@dataclass
class Plan:
    n_qubits: int
    g_bits: int
    steps: list = field(default_factory=list)      # ('local', [Pass]) | ('swap', np.int32[k,5])
    final_phys_of: list = field(default_factory=list)
    n_swaps: int = 0


def _qubits(g: Gate):
    return (g[1], g[2]) if g[0] in TWO_QUBIT else (g[1],)


def plan_sharded(n_qubits: int, g_bits: int, gates: Sequence[Gate], tile_bits: int = 12, low_bits: int = 3,
                 max_fused: int = 32, compile_local: bool = True) -> Plan:
    n_local = n_qubits - g_bits
    phys_of = list(range(n_qubits))                 # logical -> physical
    plan = Plan(n_qubits, g_bits)
    remaining = list(gates)
    cnot = 3                                        # QFX_GATE_CNOT

    def emit_local(batch):
        plan.steps.append(("local", compile_passes(n_local, batch, tile_bits, low_bits) if compile_local else batch))

    while remaining:
        n_before = len(remaining)
        blocked, batch, rest = set(), [], []
        for g in remaining:
            qs = _qubits(g)
            if not (set(qs) & blocked) and all(phys_of[q] < n_local for q in qs):
                batch.append((g[0], phys_of[g[1]], phys_of[g[2]] if g[0] in TWO_QUBIT else -1, g[3]))
            else:
                blocked |= set(qs)
                rest.append(g)
        remaining = rest
        if not remaining:
            if batch:
                emit_local(batch)
            break
        # Which local qubits leave?  The g whose next use is farthest away (Belady); they must sit in the
        # top g local positions, because only those exchange as contiguous slices.  Qubits that are needed
        # soon but happen to sit there are first moved down with a SWAP (three CNOTs in a local pass).
        next_use = {}
        for i, g in enumerate(remaining):
            for q in _qubits(g):
                next_use.setdefault(q, i)
        log_at = {phys_of[q]: q for q in range(n_qubits)}
        locals_ = [q for q in range(n_qubits) if phys_of[q] < n_local]
        top = [n_local - g_bits + i for i in range(g_bits)]
        order = sorted(locals_, key=lambda q: (-next_use.get(q, 1 << 60), -phys_of[q]))
        victims = order[:g_bits]
        movers = [v for v in victims if phys_of[v] not in top]
        relabel = []
        for t in top:
            if log_at[t] in victims:
                continue
            v = movers.pop()
            pa, pb = phys_of[v], t
            relabel += [(cnot, pa, pb, -1), (cnot, pb, pa, -1), (cnot, pa, pb, -1)]
            other = log_at[t]
            phys_of[v], phys_of[other] = t, pa
            log_at[t], log_at[pa] = v, other
        # the relabelling CNOTs ride in the same compiled batch as the gates before them: they often fit into a
        # pass that is needed anyway instead of costing a sweep over the shard of their own
        if batch or relabel:
            emit_local(batch + relabel)
        incoming = {log_at[n_local + i]: i for i in range(g_bits)}     # logical -> bit of the slice index
        blocked, fused, rest = set(), [], []
        for g in remaining:
            qs = _qubits(g)
            if not (set(qs) & blocked) and all(q in incoming for q in qs) and len(fused) < max_fused:
                fused.append((g[0], incoming[g[1]], incoming[g[2]] if g[0] in TWO_QUBIT else -1, g[3], -1))
            else:
                blocked |= set(qs)
                rest.append(g)
        plan.steps.append(("swap", np.asarray(fused, dtype=np.int32).reshape(-1, 5)))
        plan.n_swaps += 1
        for i in range(g_bits):
            a, b = log_at[n_local + i], log_at[n_local - g_bits + i]
            phys_of[a], phys_of[b] = n_local - g_bits + i, n_local + i
        remaining = rest
        if len(remaining) == n_before and not fused:
            # the swap itself is progress only if the next local scan can schedule something
            nxt = remaining[0]
            if not all(phys_of[q] < n_local for q in _qubits(nxt)):
                raise RuntimeError(f"sharded planner made no progress at gate {nxt}")
    plan.final_phys_of = phys_of
    return plan


class ShardedState:
    def __init__(self, n_qubits: int, world: int, rank: int = 0, device=None, loopback: bool = False):
        self.n, self.world, self.rank, self.loopback = n_qubits, world, rank, loopback
        self.g = int(np.log2(world))
        assert 1 << self.g == world and world in (2, 4, 8)
        self.n_local = n_qubits - self.g
        self.device = torch.device(device or "cuda")
        size = 1 << self.n_local
        self.epoch = 0
        if loopback:
            self.bufs = [[torch.zeros(size, dtype=torch.complex64, device=self.device) for _ in range(world)]
                         for _ in range(2)]
            self.tables = [torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=self.device)
                           for bufs in self.bufs]
            self.flag_table = None
        else:
            from .symm import SymmetricBuffer
            self.symm = [SymmetricBuffer(size * 8, rank, world, self.device) for _ in range(2)]
            self.bufs = [[s.view(torch.float32).view(size, 2)] for s in self.symm]
            self.bufs = [[torch.view_as_complex(b[0])] for b in self.bufs]
            self.tables = [s.peer_table() for s in self.symm]
            self.flags = SymmetricBuffer(256, rank, world, self.device)
            self.flag_table = self.flags.peer_table()
        self.cur = 0

    def _ranks(self):
        return range(self.world) if self.loopback else [self.rank]

    def shard(self, r=None) -> torch.Tensor:
        return self.bufs[self.cur][r if self.loopback else 0]

    def reset_zero_state(self):
        for r in self._ranks():
            s = self.shard(r)
            s.zero_()
            if r == 0:
                s[0] = 1.0

    def run(self, plan: Plan, theta: torch.Tensor, angles: torch.Tensor | None = None):
        th = theta.reshape(1, -1).contiguous().float()
        ang = None if angles is None else angles.reshape(1, -1).contiguous().float()
        for kind, payload in plan.steps:
            if kind == "local":
                for r in self._ranks():
                    run_passes(self.shard(r)[None], self.n_local, payload, th, ang)
            else:
                ops = torch.as_tensor(payload, device=self.device).contiguous() if len(payload) else \
                    torch.zeros(1, 5, dtype=torch.int32, device=self.device)
                self.epoch += 1
                nxt = 1 - self.cur
                for r in self._ranks():
                    dst = self.bufs[nxt][r if self.loopback else 0]
                    _lib.call("qfx_sv_global_swap_fused", _lib.ptr(self.tables[self.cur]),
                              _lib.ptr(torch.view_as_real(dst)), _lib.c_int(self.n_local), _lib.c_int(self.g),
                              _lib.c_int(r), _lib.c_int(self.world), _lib.ptr(ops), _lib.c_int(len(payload)),
                              _lib.ptr(th), _lib.c_int(th.shape[1]), _lib.ptr(ang),
                              _lib.c_int(0 if ang is None else ang.shape[-1]),
                              _lib.ptr(self.flag_table), _lib.c_u32(self.epoch), _lib.stream_ptr())
                self.cur = nxt

    def expvals(self, plan: Plan) -> torch.Tensor:
        """<Z_q> per LOGICAL qubit (partial sums of every shard added; all_reduce in multi-process mode)."""
        tot = torch.zeros(self.n, device=self.device)
        for r in self._ranks():
            tot += expvals_z(self.shard(r)[None], self.n_local, self.n, r)[0]
        if not self.loopback and self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(tot)
        phys = torch.as_tensor(plan.final_phys_of, device=self.device, dtype=torch.long)
        return tot[phys]

    def expvals_parity(self, plan: Plan, masks: Sequence[int]) -> torch.Tensor:
        """sum_i |psi_i|^2 (-1)^parity(i & masks[q]) with masks over LOGICAL qubits (ops/large.py:
        ring_readout_masks): the readout of a circuit whose final CNOT ring was not simulated."""
        from ..ops.large import expvals_parity
        phys = [sum(1 << plan.final_phys_of[l] for l in range(self.n) if (m >> l) & 1) for m in masks]
        tot = torch.zeros(len(masks), device=self.device)
        for r in self._ranks():
            tot += expvals_parity(self.shard(r)[None], self.n_local, phys, r)[0]
        if not self.loopback and self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(tot)
        return tot

    def gather_logical(self, plan: Plan) -> torch.Tensor:
        """Full state in logical qubit order (loopback / small states only; for tests)."""
        assert self.loopback
        full = torch.cat([self.shard(r) for r in range(self.world)])        # physical order, rank = top bits
        idx = torch.arange(1 << self.n, device=self.device)
        src = torch.zeros_like(idx)
        for q, p in enumerate(plan.final_phys_of):                            # logical bit q lives at physical bit p
            src |= ((idx >> q) & 1) << p
        return full[src]


def nccl_swap_baseline(shard: torch.Tensor, world: int) -> torch.Tensor:
    """B1: the global<->top-local qubit exchange as one NCCL all_to_all (no fused gates)."""
    import torch.distributed as dist
    out = torch.empty_like(shard)
    dist.all_to_all_single(torch.view_as_real(out), torch.view_as_real(shard))
    return out
