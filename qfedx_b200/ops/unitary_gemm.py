"""K2: apply one ansatz unitary to a large batch of states as a complex GEMM on the tcgen05 tensor
cores (csrc/sv_gemm_tcgen05.cu).  Used for inference / evaluation where thousands of samples share
the same parameters: build U once with the register kernel (2^n basis circuits), then one GEMM.

``ansatz_unitary`` and ``apply_unitary_reference`` are the torch reference of the same op.
"""
from __future__ import annotations

import torch

from . import _lib, ref
from .statevector import BASIS, MAX_REG_QUBITS, sv_forward

MIN_QUBITS, MAX_QUBITS = 6, 12


def ansatz_unitary_t(theta: torch.Tensor, entangler: str = "cnot") -> torch.Tensor:
    """U^T as complex64 [2^n, 2^n]: row b is U|b> - the 2^n basis states pushed through the ansatz by the register
    kernel K1 (n <= 10) or, for 11 and 12 qubits where no register kernel exists, by the pass engine K5."""
    theta = theta if theta.dim() == 4 else theta[None]
    L, n = theta.shape[1], theta.shape[2]
    if n <= MAX_REG_QUBITS:
        eye = torch.eye(1 << n, device=theta.device, dtype=torch.float32)
        _, psi = sv_forward(None, theta, None, "ry", entangler, init_state=eye, want_state=True)
        return psi
    from .large import ansatz_gates, compile_passes, run_passes
    psi = torch.eye(1 << n, device=theta.device, dtype=torch.complex64)
    passes = compile_passes(n, ansatz_gates(n, L, entangler, feature_map=False))
    run_passes(psi, n, passes, theta.reshape(1, -1).contiguous().float(), None)
    return psi


def prefers_unitary_gemm(n_qubits: int, n_layers: int, batch: int) -> bool:
    """Where K2 is the faster inference engine (measured on B200, bench aux `unitary_gemm_vs_engines`): its cost does not
    depend on the circuit depth - one [2 x 2^n]^2 GEMM per batch - while the register kernel (n <= 10) and the pass engine
    (n > 10) grow linearly with it.  Break-even depths: about 12 layers at 10 and 12 qubits, 8 at 11, 10 at 8-9 qubits."""
    if not (MIN_QUBITS <= n_qubits <= MAX_QUBITS):
        return False
    if n_qubits > MAX_REG_QUBITS:
        return batch >= 256 and n_layers >= (12 if n_qubits >= 12 else 8)
    return batch >= 4096 and n_qubits >= 8 and n_layers >= (12 if n_qubits >= 10 else 10)


class UnitaryGemm:
    """Holds the packed unitary of one parameter set; ``expvals(angles)`` / ``apply(states)``."""

    def __init__(self, theta: torch.Tensor, entangler: str = "cnot", basis: str = "ry", precision: int = 3):
        theta = theta.detach()
        theta = theta if theta.dim() == 4 else theta[None]
        self.n = theta.shape[2]
        if not (MIN_QUBITS <= self.n <= MAX_QUBITS):
            raise _lib.QfxError(f"unitary GEMM supports {MIN_QUBITS}..{MAX_QUBITS} qubits, got {self.n}")
        if not theta.is_cuda:
            raise _lib.QfxError("UnitaryGemm needs CUDA tensors")
        self.basis, self.precision, self.device = basis, precision, theta.device
        self.m = 2 << self.n
        ut = ansatz_unitary_t(theta, entangler).contiguous()
        self.a_hi = torch.empty(self.m, self.m, device=self.device)
        self.a_lo = torch.empty_like(self.a_hi)
        _lib.call("qfx_unitary_pack", _lib.ptr(torch.view_as_real(ut)), _lib.c_int(self.n), _lib.ptr(self.a_hi),
                  _lib.ptr(self.a_lo), _lib.stream_ptr())

    def _pad(self, batch: int) -> int:
        return (batch + 127) // 128 * 128 if batch > 64 else 64

    def _gemm(self, b_hi, b_lo, bp, want_states, want_expvals):
        out = torch.empty(bp, self.m, device=self.device) if want_states else None
        ez = torch.empty(bp, self.n, device=self.device) if want_expvals else None
        _lib.call("qfx_unitary_gemm", _lib.ptr(self.a_hi), _lib.ptr(self.a_lo), _lib.ptr(b_hi), _lib.ptr(b_lo),
                  _lib.c_int(self.n), _lib.c_int(bp), _lib.c_int(self.precision), _lib.ptr(out), _lib.ptr(ez),
                  _lib.stream_ptr())
        return out, ez

    def expvals(self, angles: torch.Tensor, want_states: bool = False):
        """<Z_q> [B, n] for feature-map angles [B, n]; the input states never touch HBM as complex
        tensors: they are generated straight into the GEMM operand layout."""
        B = angles.shape[0]
        bp = self._pad(B)
        b_hi = torch.empty(bp, self.m, device=self.device)
        b_lo = torch.empty_like(b_hi)
        _lib.call("qfx_unitary_pack_product_states", _lib.ptr(angles.contiguous().float()), _lib.c_int(self.n),
                  _lib.c_int(BASIS[self.basis]), _lib.c_int(B), _lib.c_int(bp), _lib.ptr(b_hi), _lib.ptr(b_lo),
                  _lib.stream_ptr())
        out, ez = self._gemm(b_hi, b_lo, bp, want_states, True)
        if want_states:
            dim = 1 << self.n
            return ez[:B], torch.complex(out[:B, :dim], out[:B, dim:])
        return ez[:B]

    def apply(self, states: torch.Tensor) -> torch.Tensor:
        """complex64 [B, 2^n] -> U states."""
        B, dim = states.shape
        bp = self._pad(B)
        flat = torch.cat([states.real, states.imag], dim=1).contiguous().float()
        b_hi = torch.empty(bp, self.m, device=self.device)
        b_lo = torch.empty_like(b_hi)
        _lib.call("qfx_unitary_split_states", _lib.ptr(flat), _lib.c_int(self.n), _lib.c_int(B), _lib.c_int(bp),
                  _lib.ptr(b_hi), _lib.ptr(b_lo), _lib.stream_ptr())
        out, _ = self._gemm(b_hi, b_lo, bp, True, False)
        return torch.complex(out[:B, :dim], out[:B, dim:])


def apply_unitary_reference(theta: torch.Tensor, states: torch.Tensor, entangler: str = "cnot") -> torch.Tensor:
    """torch reference: dense U (ops/ref.py) times the states."""
    th = theta if theta.dim() == 3 else theta[0]
    U = ref.ansatz_unitary_ref(th.double().cpu(), entangler)
    return (U @ states.cpu().to(torch.complex128).T).T
