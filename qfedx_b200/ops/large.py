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


def ansatz_gates(n: int, n_layers: int, entangler: str = "cnot", feature_map: bool = True,
                 final_ring: bool = True) -> List[Tuple[int, int, int, int]]:
    """(kind, q0, q1, param) list of the VQC: RY feature map + L x [RX RZ per qubit, ring].  With
    ``final_ring=False`` the last layer's ring is left out (read <Z> through ``ring_readout_masks``)."""
    gates = []
    if feature_map:
        gates += [(GATE["ry_feature"], q, -1, q) for q in range(n)]
    for l in range(n_layers):
        for q in range(n):
            gates.append((GATE["rx"], q, -1, (l * n + q) * 2))
            gates.append((GATE["rz"], q, -1, (l * n + q) * 2 + 1))
        if entangler != "none" and (final_ring or l < n_layers - 1):
            gates += [(GATE[entangler], c, t, -1) for c, t in ring_pairs(n)]
    return gates


def ring_readout_masks(n: int, entangler: str = "cnot") -> List[int]:
    """Bit masks m_q with <Z_q>(ring |psi>) = sum_i |psi_i|^2 (-1)^parity(i & m_q).

    A CNOT ring permutes basis states; in the Heisenberg picture CNOT(c,t) maps Z_t -> Z_c Z_t and leaves
    Z_c alone, so pulling Z_q back through the ring (last gate first) gives a Z-string.  CZ is diagonal: the
    masks stay single bits.  The readout of a VQC therefore never needs the final ring to be simulated."""
    masks = [1 << q for q in range(n)]
    if entangler == "cnot":
        for c, t in reversed(ring_pairs(n)):
            masks = [m ^ (1 << c) if (m >> t) & 1 else m for m in masks]
    return masks


@dataclass
class Pass:
    tile_qubits: np.ndarray   # int32 [T] host
    ops: np.ndarray           # int32 [k,5] host
    ops_dev: torch.Tensor | None = None
    plan: np.ndarray | None = None       # segment plan for the register-resident kernel (12-qubit tiles)
    plan_dev: torch.Tensor | None = None


REG_TILE_BITS = 12
MAX_OPS_PER_PASS = 160        # csrc/sv_large.cu: MAX_OPS_PER_PASS
PLAN_KIND_RXRZ_STRIPPED = 7   # plan-private op kind (csrc/sv_large.cu: V2_KIND_RXRZ_STRIPPED)


def plan_segments(ops: np.ndarray) -> np.ndarray:
    """Split a 12-bit pass into segments for the register-resident kernel (csrc/sv_large.cu, v2).

    Positions: 0-4 lane bits, 5-8 register bits, 9-11 warp bits (inactive in a segment).  Arithmetic on a
    register bit is plain FMAs; on a lane bit every amplitude costs two warp shuffles more, so within a segment
    one-qubit gates and CNOT targets are only executed on REGISTER bits - anything else waits for a later
    segment whose register set holds its bit (controls and CZ operands may sit on lane bits).  Each segment
    picks, among the bits whose gates are due next, the register set that lets the most work run.

    The tile lives in shared memory with its low index nibble XOR-ed with the two higher nibbles, so a segment's
    re-layout is bank-conflict free whenever four of the five lane bits fall into distinct classes mod 4; the
    planner keeps one bit of every class out of the register set for that.

    Returns the int32 plan: n_seg | 1 << 16, then per segment pos_to_bit[12], n_ops, n_ops x (kind, posA, posB,
    op index), then the trailer described below."""
    from itertools import combinations
    T = REG_TILE_BITS
    cnot, cz = GATE["cnot"], GATE["cz"]
    kinds = [int(o[0]) for o in ops]
    b0s, b1s = [int(o[1]) for o in ops], [int(o[2]) for o in ops]

    def bits_of(i):
        return (b0s[i], b1s[i]) if kinds[i] in TWO_QUBIT else (b0s[i],)

    def reg_bits(i):        # bits that must be register bits for op i to run
        return () if kinds[i] == cz else ((b1s[i],) if kinds[i] == cnot else (b0s[i],))

    def run(remaining, regs, active):
        blocked, chosen, rest, score = set(), [], [], 0
        for i in remaining:
            bs = bits_of(i)
            if any(b in blocked or b not in active for b in bs) or any(b not in regs for b in reg_bits(i)):
                blocked.update(bs)
                rest.append(i)
            else:
                chosen.append(i)
                score += 1 if kinds[i] in TWO_QUBIT else 3
        return chosen, rest, score

    remaining = list(range(len(ops)))
    segs = []
    while remaining:
        next_use, due = {}, []                     # first use of every bit; bits in the order their gates fall due
        for rank, i in enumerate(remaining):
            for b in bits_of(i):
                next_use.setdefault(b, rank)
            for b in reg_bits(i):
                if b not in due:
                    due.append(b)
        first = remaining[0]
        must_reg, must_active = set(reg_bits(first)), set(bits_of(first))
        pool = due[:7] + [b for b in sorted(range(T), key=lambda b: next_use.get(b, 1 << 30)) if b not in due[:7]]
        pool = pool[:max(7, 4)]
        best = None
        for regs in combinations(pool, 4):
            regs = set(regs)
            if not must_reg <= regs or any(all(b in regs for b in range(c, T, 4)) for c in range(4)):
                continue
            free = [b for b in range(T) if b not in regs]
            by_need = sorted(free, key=lambda b: (b not in must_active, next_use.get(b, 1 << 30)))
            lanes = []
            for b in by_need:                      # one lane bit per class mod 4, most urgent first
                if all(b % 4 != l % 4 for l in lanes):
                    lanes.append(b)
            lanes = lanes[:4]
            lanes.append(next(b for b in by_need if b not in lanes))
            if not must_active <= regs | set(lanes):
                continue
            chosen, rest, score = run(remaining, regs, regs | set(lanes))
            if best is None or score > best[0]:
                best = (score, sorted(regs), lanes, chosen, rest)
        assert best is not None and best[3], "segment planner made no progress"
        _, regs, lanes, chosen, rest = best
        warps = sorted(b for b in range(T) if b not in regs and b not in lanes)
        p2b = lanes + regs + warps
        pos_of = {b: p for p, b in enumerate(p2b)}
        seg = list(p2b) + [len(chosen)]
        for i in chosen:
            seg += [kinds[i], pos_of[b0s[i]], pos_of[b1s[i]] if kinds[i] in TWO_QUBIT else -1, i]
        segs.append(seg)
        remaining = rest
    # RXRZ gates on register bits run without their scalar phase e^{-i phi/2} (12 instead of 16 instructions per
    # amplitude pair); the kernel folds the product of those phases into one carrier gate of the pass, which keeps
    # its full matrix.  Trailer: mode per op (0 general, 1 stripped, 2 carrier), then the carrier's op index.
    modes, carrier = [0] * len(ops), -1
    for sgm in segs:
        for e in range(13, len(sgm), 4):
            kind, pos_a, op_i = sgm[e], sgm[e + 1], sgm[e + 3]
            if kind == GATE["rxrz"] and pos_a >= 5:
                if carrier < 0:
                    carrier, modes[op_i] = op_i, 2
                else:
                    modes[op_i], sgm[e] = 1, PLAN_KIND_RXRZ_STRIPPED
    flat = [len(segs) | (1 << 16)]
    for sgm in segs:
        flat += sgm
    flat += modes + [carrier]
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
    # a deep circuit on few qubits can put more gates into one tile than a launch takes (MAX_OPS_PER_PASS in
    # csrc/sv_large.cu): such a pass is cut into consecutive passes over the same tile
    passes = []
    for i in range(n_pass):
        po = ops[begin[i]:begin[i + 1]]
        for lo in range(0, max(1, len(po)), MAX_OPS_PER_PASS):
            passes.append(Pass(tq[i].copy(), po[lo:lo + MAX_OPS_PER_PASS].copy()))
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


def expvals_parity(states: torch.Tensor, n_local: int, masks: Sequence[int], high_bits: int = 0) -> torch.Tensor:
    """out[b, q] = sum_i |psi_i|^2 (-1)^parity((high_bits << n_local | i) & masks[q]); states [B, 2^n_local]."""
    B, m = states.shape[0], np.asarray(masks, dtype=np.uint64)
    out = torch.empty(B, len(m), device=states.device, dtype=torch.float32)
    _lib.call("qfx_sv_large_expz_masks", _lib.ptr(torch.view_as_real(states)), _lib.c_int(n_local), _lib.c_int(len(m)),
              _lib.c_u32(high_bits), m.ctypes.data_as(ctypes.c_void_p), _lib.c_int(B), _lib.ptr(out), _lib.stream_ptr())
    return out


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
        # readout path: the final ring is folded into parity masks instead of being simulated
        self.readout_gates = ansatz_gates(n_qubits, n_layers, entangler, final_ring=False)
        self.readout_masks = ring_readout_masks(n_qubits, entangler if n_layers > 0 else "none")
        self.readout_passes = (compile_passes(n_qubits, self.readout_gates, self.tile_bits, self.low_bits)
                               if _lib.available() else None)

    def states(self, angles: torch.Tensor, theta: torch.Tensor, shift_param=None, shift_value=None,
               batch: int | None = None, before_final_ring: bool = False) -> torch.Tensor:
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
            return ref.vqc_state_ref(angles, th, None if th.shape[0] == 1 else torch.arange(B) % th.shape[0], "ry", self.entangler)
        st = torch.empty(B, 1 << self.n, dtype=torch.complex64, device=angles.device)
        run_passes(st, self.n, self.readout_passes if before_final_ring else self.passes, theta, angles, shift_param,
                   shift_value, init_zero=True, low_bits=self.low_bits, use_register_engine=self.register_engine)
        return st

    def _readout(self, angles, theta, shift_param=None, shift_value=None, batch=None, n_out=None) -> torch.Tensor:
        """<Z_q>, q < n_out (default: every qubit), of the full circuit for every state of the batch."""
        n_out = n_out or self.n
        if not angles.is_cuda:
            return ref.expval_z_all(self.states(angles, theta, shift_param, shift_value, batch)).float()[:, :n_out]
        st = self.states(angles, theta, shift_param, shift_value, batch, before_final_ring=True)
        return expvals_parity(st, self.n, self.readout_masks[:n_out])

    def expvals(self, angles: torch.Tensor, theta: torch.Tensor, n_out: int | None = None) -> torch.Tensor:
        """A classifier reads only its first n_logits qubits: pass ``n_out`` and the readout kernel stays
        bandwidth-bound instead of evaluating n parities per amplitude."""
        return self._readout(angles, theta, n_out=n_out)

    def param_shift_grad(self, angles: torch.Tensor, theta: torch.Tensor, cot: torch.Tensor,
                         max_states: int = 256) -> torch.Tensor:
        """d sum_s cot[s] . <Z>(angles[s]) / d theta by the two-term shift rule; theta is [P]; cot is
        [S, n_out] - only the first n_out = cot.shape[1] qubits are read out."""
        n_out = cot.shape[1]
        S, P, dev = angles.shape[0], self.n_params, angles.device
        grad = torch.zeros(P, device=dev)
        per_chunk = max(1, max_states // (2 * S))          # parameters per chunk
        # The last layer's RZ angles cannot move a Z-basis readout: RZ is diagonal and what follows it (the final
        # ring) only permutes or re-phases basis states, so |psi_i|^2 is unchanged - their gradient is exactly 0
        # and their 2n shifted circuits are not simulated.
        last_rz = {(self.L - 1) * self.n * 2 + 2 * q + 1 for q in range(self.n)} if self.L > 0 else set()
        live = [p for p in range(P) if p not in last_rz]
        # A shifted angle that first appears in pass k leaves passes 0..k-1 untouched: those run once per
        # minibatch on the unshifted circuit and every shifted state of the group starts from a copy.
        groups = {}
        for p in live:
            groups.setdefault(self._first_pass_of_param().get(p, 0) if angles.is_cuda else 0, []).append(p)
        th = theta.reshape(1, P)
        prefix, prefix_k = None, 0
        for k in sorted(groups):
            if k > 0:
                if prefix is None:
                    prefix = torch.empty(S, 1 << self.n, dtype=torch.complex64, device=dev)
                run_passes(prefix, self.n, self.readout_passes[prefix_k:k], th, angles, init_zero=(prefix_k == 0),
                           low_bits=self.low_bits, use_register_engine=self.register_engine)
                prefix_k = k
            for c0 in range(0, len(groups[k]), per_chunk):
                ps = torch.tensor(groups[k][c0:c0 + per_chunk], device=dev)
                # item order (param, sign, sample): row -> sample = row % S keeps angle_rows = S valid
                sp = ps.repeat_interleave(2 * S).to(torch.int32)
                sv = torch.tensor([np.pi / 2, -np.pi / 2], device=dev).repeat_interleave(S).repeat(len(ps)).float()
                if k == 0:
                    E = self._readout(angles, th, sp, sv, batch=len(sp), n_out=n_out)
                else:
                    st = prefix.repeat(2 * len(ps), 1)
                    run_passes(st, self.n, self.readout_passes[k:], th, angles, sp, sv, init_zero=False,
                               low_bits=self.low_bits, use_register_engine=self.register_engine)
                    E = expvals_parity(st, self.n, self.readout_masks[:n_out])
                val = (E.reshape(len(ps), 2, S, n_out) * cot[None, None]).sum(dim=(2, 3))
                grad[ps] = 0.5 * (val[:, 0] - val[:, 1])
        return grad

    def adjoint_grad(self, angles: torch.Tensor, theta: torch.Tensor, cot: torch.Tensor) -> torch.Tensor:
        """Same gradient as ``param_shift_grad`` by adjoint differentiation: one forward sweep, lambda = O psi, one
        backward sweep over the compiled passes carrying psi and lambda (csrc/sv_large.cu: sv_large_adjoint_pass_kernel)
        - a constant number of state sweeps instead of 2P shifted circuits.

        theta [P]: shared by all S samples, returns [P].  theta [R, P]: sample b runs with row b % R (several clients'
        models in one batch, rows interleaved) and the result is [R, P], row r summed over the samples of row r."""
        n_out = cot.shape[1]
        S, P, dev = angles.shape[0], self.n_params, angles.device
        shared = theta.dim() == 1
        th = theta.reshape(-1, P).contiguous().float()
        R = th.shape[0]
        if not angles.is_cuda:
            t = th.detach().clone().requires_grad_(True)
            gid = None if R == 1 else (torch.arange(S) % R)
            E = ref.expval_z_all(ref.vqc_state_ref(angles, t.reshape(R, self.L, self.n, 2), gid, "ry", self.entangler)).float()
            g = torch.autograd.grad((E[:, :n_out] * cot).sum(), t)[0]
            if self.L > 0:   # exactly-zero gradients of the last layer's RZ angles (see below): no rounding noise for Adam
                g[:, [(self.L - 1) * self.n * 2 + 2 * q + 1 for q in range(self.n)]] = 0.0
            return g.reshape(P) if shared else g
        ang = angles.contiguous().float()
        psi = self.states(ang, th, before_final_ring=True)
        lam = torch.empty_like(psi)
        masks = (ctypes.c_ulonglong * n_out)(*[int(m) for m in self.readout_masks[:n_out]])
        _lib.call("qfx_sv_large_parity_obs", _lib.ptr(torch.view_as_real(psi)), _lib.ptr(torch.view_as_real(lam)),
                  _lib.c_int(self.n), _lib.c_int(n_out), masks, _lib.ptr(cot.contiguous().float()), _lib.c_int(S),
                  _lib.stream_ptr())
        grad = torch.zeros(R, P, device=dev)
        for p in reversed(self.readout_passes):
            if p.ops_dev is None or p.ops_dev.device != dev:
                p.ops_dev = torch.as_tensor(p.ops, device=dev).contiguous()
            T = len(p.tile_qubits)
            tq = np.ascontiguousarray(p.tile_qubits, dtype=np.int32)
            low = 0
            while low < min(self.low_bits, T) and tq[low] == low:
                low += 1
            _lib.call("qfx_sv_large_adjoint_pass", _lib.ptr(torch.view_as_real(psi)), _lib.ptr(torch.view_as_real(lam)),
                      _lib.c_int(self.n), _lib.c_int(S), _lib.c_int(T), _lib.c_int(low), tq.ctypes.data_as(ctypes.c_void_p),
                      _lib.ptr(p.ops_dev), _lib.c_int(len(p.ops)), _lib.ptr(th), _lib.c_int(R), _lib.c_int(P), _lib.ptr(ang),
                      _lib.c_int(S), _lib.c_int(ang.shape[1]), _lib.ptr(grad), _lib.stream_ptr())
        if self.L > 0:
            # the last layer's RZ angles cannot move a Z-basis readout (see param_shift_grad): their gradient is exactly 0;
            # the sweep returns rounding noise there, which Adam would normalise into full-size steps
            last_rz = torch.tensor([(self.L - 1) * self.n * 2 + 2 * q + 1 for q in range(self.n)], device=dev)
            grad[:, last_rz] = 0.0
        return grad.reshape(P) if shared else grad

    def _first_pass_of_param(self) -> dict:
        """parameter index -> index of the (only) readout pass whose gates use it."""
        if getattr(self, "_param_pass", None) is None:
            self._param_pass = {}
            for k, p in enumerate(self.readout_passes):
                for op in p.ops:
                    kind = int(op[0])
                    if kind in (GATE["rx"], GATE["ry"], GATE["rz"], GATE["rxrz"]):
                        self._param_pass.setdefault(int(op[3]), k)
                    if kind == GATE["rxrz"]:
                        self._param_pass.setdefault(int(op[4]), k)
        return self._param_pass
