"""Large-state VQC engine (K5): compiled multi-gate passes over statevectors in global memory, batched
over states (clients or parameter-shift terms).  csrc/circuit_passes.cu schedules, csrc/sv_large.cu runs.
CPU tensors use the torch reference (ops/ref.py), which is also the numerics oracle in the tests."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Sequence, Tuple

import numpy as np
import torch

from . import _lib, ref
from ..oracle.statevector import ring_pairs

GATE = {"rx": 0, "ry": 1, "rz": 2, "cnot": 3, "cz": 4, "rxrz": 5, "ry_feature": 6}
TWO_QUBIT = (GATE["cnot"], GATE["cz"])


def ansatz_gates(n: int, n_layers: int, entangler: str = "cnot", feature_map: bool = True) -> List[Tuple[int, int, int, int]]:
    """(kind, q0, q1, param) list of the VQC: RY feature map + L x [RX RZ per qubit, ring]."""
    gates = []
    if feature_map:
        gates += [(GATE["ry_feature"], q, -1, q) for q in range(n)]
    for l in range(n_layers):
        for q in range(n):
            gates.append((GATE["rx"], q, -1, (l * n + q) * 2))
            gates.append((GATE["rz"], q, -1, (l * n + q) * 2 + 1))
        if entangler != "none":
            gates += [(GATE[entangler], c, t, -1) for c, t in ring_pairs(n)]
    return gates


@dataclass
class Pass:
    tile_qubits: np.ndarray   # int32 [T] host
    ops: np.ndarray           # int32 [k,5] host
    ops_dev: torch.Tensor | None = None
    plan: np.ndarray | None = None       # segment plan for the register-resident kernel (12-qubit tiles)
    plan_dev: torch.Tensor | None = None


REG_TILE_BITS = 12


def plan_segments(ops: np.ndarray) -> np.ndarray:
    """Split a 12-bit pass into segments for the register-resident kernel (csrc/sv_large.cu, v2).

    Positions: 0-4 lane bits, 5-8 register bits, 9-11 warp bits (inactive in a segment).  Tile bits 0-3
    stay on lane positions 0-3 (conflict-free shared-memory re-layout); the other eight bits are
    re-assigned per segment: the three whose next use is farthest become inactive (Belady), the busiest
    four of the rest become register bits.  Returns the int32 plan: n_seg, then per segment
    pos_to_bit[12], n_ops, n_ops x (kind, posA, posB, op index)."""
    flexible = list(range(4, REG_TILE_BITS))
    remaining = list(range(len(ops)))

    def bits_of(i):
        k, b0, b1 = int(ops[i][0]), int(ops[i][1]), int(ops[i][2])
        return (b0, b1) if k in TWO_QUBIT else (b0,)

    segs = []
    while remaining:
        next_use = {}
        for rank, i in enumerate(remaining):
            for b in bits_of(i):
                next_use.setdefault(b, rank)
        order = sorted(flexible, key=lambda b: -next_use.get(b, 1 << 30))
        inactive = set(order[:3])
        blocked, chosen, rest = set(), [], []
        for i in remaining:
            bs = set(bits_of(i))
            if bs & inactive or bs & blocked:
                blocked |= bs
                rest.append(i)
            else:
                chosen.append(i)
        assert chosen, "segment planner made no progress"
        active = [b for b in flexible if b not in inactive]          # 5 bits: 4 register + 1 lane
        use = {b: 0 for b in active}
        for i in chosen:
            for b in bits_of(i):
                if b in use:
                    use[b] += 1 if int(ops[i][0]) in TWO_QUBIT else 2      # 1q gates gain most from registers
        by_use = sorted(active, key=lambda b: -use[b])
        regs, lane4 = sorted(by_use[:4]), by_use[4]
        p2b = [0, 1, 2, 3, lane4] + regs + sorted(inactive)
        pos_of = {b: p for p, b in enumerate(p2b)}
        seg = list(p2b) + [len(chosen)]
        for i in chosen:
            k, b0, b1 = int(ops[i][0]), int(ops[i][1]), int(ops[i][2])
            seg += [k, pos_of[b0], pos_of[b1] if k in TWO_QUBIT else -1, i]
        segs.append(seg)
        remaining = rest
    flat = [len(segs)]
    for sgm in segs:
        flat += sgm
    return np.asarray(flat, dtype=np.int32)


def compile_passes(n_qubits: int, gates: Sequence[Tuple[int, int, int, int]], tile_bits: int = 12,
                   low_bits: int = 3) -> List[Pass]:
    """Run the C++ pass compiler (host code in the CUDA library; needs no GPU)."""
    tile_bits, low_bits = min(tile_bits, n_qubits), min(low_bits, tile_bits, n_qubits)
    g = np.ascontiguousarray(np.asarray(gates, dtype=np.int32).reshape(-1, 4))
    max_passes, max_ops = max(4, len(g)), max(4, len(g))
    tq = np.zeros((max_passes, tile_bits), dtype=np.int32)
    begin = np.zeros(max_passes + 1, dtype=np.int32)
    ops = np.zeros((max_ops, 5), dtype=np.int32)
    ip = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    fn = _lib.lib().qfx_circuit_compile
    fn.restype = ctypes.c_int
    n_pass = fn(_lib.c_int(n_qubits), _lib.c_int(len(g)), ip(g), _lib.c_int(tile_bits), _lib.c_int(low_bits),
                _lib.c_int(max_passes), _lib.c_int(max_ops), ip(tq), ip(begin), ip(ops))
    if n_pass < 0:
        raise _lib.QfxError(f"qfx_circuit_compile failed ({n_pass}): {_lib.lib().qfx_last_error().decode()}")
    passes = [Pass(tq[i].copy(), ops[begin[i]:begin[i + 1]].copy()) for i in range(n_pass)]
    if tile_bits == REG_TILE_BITS:
        for p in passes:
            p.plan = plan_segments(p.ops)
    return passes


def run_passes(states: torch.Tensor, n_qubits: int, passes: List[Pass], theta=None, angles=None, shift_param=None,
               shift_value=None, init_zero: bool = False, low_bits: int = 3, use_register_engine: bool = True) -> None:
    """In place on ``states`` [B, 2^n] complex64 (CUDA)."""
    B = states.shape[0]
    flat = torch.view_as_real(states)
    th = None if theta is None else theta.reshape(-1, theta.shape[-1]).contiguous().float()
    ang = None if angles is None else angles.contiguous().float()
    for i, p in enumerate(passes):
        if p.ops_dev is None or p.ops_dev.device != states.device:
            p.ops_dev = torch.as_tensor(p.ops, device=states.device).contiguous()
        T = len(p.tile_qubits)
        if p.plan is not None and (p.plan_dev is None or p.plan_dev.device != states.device):
            p.plan_dev = torch.as_tensor(p.plan, device=states.device).contiguous()
        tq = np.ascontiguousarray(p.tile_qubits, dtype=np.int32)
        low = 0                                   # leading tile qubits that are 0,1,2,.. move as contiguous runs
        while low < min(low_bits, T) and tq[low] == low:
            low += 1
        _lib.call("qfx_sv_large_pass", _lib.ptr(flat), _lib.c_int(n_qubits), _lib.c_int(B), _lib.c_int(T),
                  _lib.c_int(low), tq.ctypes.data_as(ctypes.c_void_p), _lib.ptr(p.ops_dev),
                  _lib.c_int(len(p.ops)), _lib.ptr(th), _lib.c_int(0 if th is None else th.shape[0]),
                  _lib.c_int(0 if th is None else th.shape[1]), _lib.ptr(ang),
                  _lib.c_int(0 if ang is None else ang.shape[0]), _lib.c_int(0 if ang is None else ang.shape[1]),
                  _lib.ptr(shift_param), _lib.ptr(shift_value), _lib.c_int(1 if (init_zero and i == 0) else 0),
                  _lib.ptr(p.plan_dev if use_register_engine else None),
                  _lib.c_int(0 if p.plan is None else len(p.plan)), _lib.stream_ptr())


def expvals_z(states: torch.Tensor, n_local: int, n_total: int | None = None, high_bits: int = 0) -> torch.Tensor:
    n_total = n_total or n_local
    B = states.shape[0]
    out = torch.empty(B, n_total, device=states.device, dtype=torch.float32)
    _lib.call("qfx_sv_large_expz", _lib.ptr(torch.view_as_real(states)), _lib.c_int(n_local), _lib.c_int(n_total),
              _lib.c_u32(high_bits), _lib.c_int(B), _lib.ptr(out), _lib.stream_ptr())
    return out


class LargeVQC:
    """n-qubit VQC on the pass engine.  ``expvals`` batches over samples, ``param_shift_grad`` batches the
    2P shifted circuits of every sample (K4b for large states)."""

    def __init__(self, n_qubits: int, n_layers: int, entangler: str = "cnot", tile_bits: int = 12, low_bits: int = 3,
                 register_engine: bool = True):
        self.n, self.L, self.entangler = n_qubits, n_layers, entangler
        self.register_engine = register_engine
        self.tile_bits, self.low_bits = min(tile_bits, n_qubits), min(low_bits, n_qubits)
        self.gates = ansatz_gates(n_qubits, n_layers, entangler)
        self.n_params = 2 * n_qubits * n_layers
        self.passes = compile_passes(n_qubits, self.gates, self.tile_bits, self.low_bits) if _lib.available() else None
        self.n_gates = len(self.gates)

    def states(self, angles: torch.Tensor, theta: torch.Tensor, shift_param=None, shift_value=None,
               batch: int | None = None) -> torch.Tensor:
        B = batch or angles.shape[0]
        if not angles.is_cuda:
            th = theta.reshape(-1, self.L, self.n, 2)
            if shift_param is not None:
                th = th.reshape(-1, self.n_params).repeat(B // th.shape[0], 1).clone()
                rows = torch.arange(B)[shift_param >= 0]
                th[rows, shift_param[shift_param >= 0].long()] += shift_value[shift_param >= 0]
                th = th.reshape(B, self.L, self.n, 2)
                ang = angles.repeat(B // angles.shape[0], 1)
                return ref.vqc_state_ref(ang, th, torch.arange(B), "ry", self.entangler)
            return ref.vqc_state_ref(angles, th, None if th.shape[0] == 1 else torch.arange(B), "ry", self.entangler)
        st = torch.empty(B, 1 << self.n, dtype=torch.complex64, device=angles.device)
        run_passes(st, self.n, self.passes, theta, angles, shift_param, shift_value, init_zero=True,
                   low_bits=self.low_bits, use_register_engine=self.register_engine)
        return st

    def expvals(self, angles: torch.Tensor, theta: torch.Tensor) -> torch.Tensor:
        st = self.states(angles, theta)
        return expvals_z(st, self.n) if st.is_cuda else ref.expval_z_all(st).float()

    def param_shift_grad(self, angles: torch.Tensor, theta: torch.Tensor, cot: torch.Tensor,
                         max_states: int = 256) -> torch.Tensor:
        """d sum_s cot[s] . <Z>(angles[s]) / d theta by the two-term shift rule; theta is [P]."""
        S, P, dev = angles.shape[0], self.n_params, angles.device
        grad = torch.zeros(P, device=dev)
        per_chunk = max(1, max_states // (2 * S))          # parameters per chunk
        for p0 in range(0, P, per_chunk):
            ps = torch.arange(p0, min(P, p0 + per_chunk), device=dev)
            # item order (param, sign, sample): row -> sample = row % S keeps angle_rows = S valid
            sp = ps.repeat_interleave(2 * S).to(torch.int32)
            sv = torch.tensor([np.pi / 2, -np.pi / 2], device=dev).repeat_interleave(S).repeat(len(ps)).float()
            st = self.states(angles, theta.reshape(1, P), sp, sv, batch=len(sp))
            E = expvals_z(st, self.n) if st.is_cuda else ref.expval_z_all(st).float()
            val = (E.reshape(len(ps), 2, S, self.n) * cot[None, None]).sum(dim=(2, 3))
            grad[ps] = 0.5 * (val[:, 0] - val[:, 1])
        return grad
