"""Plain-PyTorch complex-tensor statevector VQC.

Two jobs: (1) the CPU fallback behind every op in ``qfedx_b200.ops`` so the framework
and its non-GPU tests run anywhere, and (2) baseline B1 of SURVEY.md section 2.5 - the
"reference-style" torch simulator the hand-written sm_100a engine is measured against
(the reference itself has no simulator beyond one 4-qubit ``Statevector.from_instruction``
call, /root/reference/src/QFed/qAmplitude.py:44-46).  Differentiable by torch autograd.
Gate conventions: see qfedx_b200/oracle/statevector.py.
"""
from __future__ import annotations

import torch

from ..oracle.statevector import ring_pairs

__all__ = ["product_state", "apply_rot", "apply_cnot", "apply_cz", "expval_z_all", "vqc_expvals_ref",
           "vqc_state_ref", "ansatz_unitary_ref"]


def _cdtype(real_dtype):
    return torch.complex128 if real_dtype == torch.float64 else torch.complex64


def product_state(angles: torch.Tensor, basis: str = "ry") -> torch.Tensor:
    """[B, n] rotation angles -> [B, 2^n] product state (little-endian)."""
    B, n = angles.shape
    c, s = torch.cos(angles / 2), torch.sin(angles / 2)
    cd = _cdtype(angles.dtype)
    if basis == "ry":
        a0, a1 = c.to(cd), s.to(cd)
    elif basis == "rx":
        a0, a1 = c.to(cd), torch.complex(torch.zeros_like(s), -s)
    elif basis == "rz":
        a0, a1 = torch.complex(c, -s), torch.zeros_like(c).to(cd)
    else:
        raise ValueError(basis)
    psi = torch.ones(B, 1, dtype=cd, device=angles.device)
    for q in range(n):  # qubit q becomes the next more-significant bit
        pair = torch.stack([a0[:, q], a1[:, q]], dim=1)  # [B, 2]
        psi = (pair[:, :, None] * psi[:, None, :]).reshape(B, -1)
    return psi


def apply_rot(psi: torch.Tensor, kind: str, angle: torch.Tensor, q: int) -> torch.Tensor:
    """exp(-i angle P/2) on qubit q; ``angle`` is [B] (one per state) or scalar."""
    B, dim = psi.shape
    v = psi.reshape(B, dim >> (q + 1), 2, 1 << q)
    angle = angle.reshape(-1, 1, 1) if angle.dim() else angle
    c, s = torch.cos(angle / 2), torch.sin(angle / 2)
    v0, v1 = v[:, :, 0], v[:, :, 1]
    if kind == "rx":
        js = torch.complex(torch.zeros_like(s), -s)
        n0, n1 = c * v0 + js * v1, js * v0 + c * v1
    elif kind == "ry":
        n0, n1 = c * v0 - s * v1, s * v0 + c * v1
    elif kind == "rz":
        ph = torch.complex(c, -s)
        n0, n1 = ph * v0, ph.conj() * v1
    else:
        raise ValueError(kind)
    return torch.stack([n0, n1], dim=2).reshape(B, dim)


def _perm_index(n, control, target, device):
    idx = torch.arange(1 << n, device=device)
    return torch.where(((idx >> control) & 1).bool(), idx ^ (1 << target), idx)


def apply_cnot(psi, control, target):
    n = psi.shape[1].bit_length() - 1
    return psi[:, _perm_index(n, control, target, psi.device)]


def apply_cz(psi, a, b):
    idx = torch.arange(psi.shape[1], device=psi.device)
    sign = 1.0 - 2.0 * (((idx >> a) & 1) & ((idx >> b) & 1)).to(psi.real.dtype)
    return psi * sign


def expval_z_all(psi: torch.Tensor) -> torch.Tensor:
    """[B, 2^n] -> [B, n] with <Z_q> for every qubit."""
    n = psi.shape[1].bit_length() - 1
    prob = psi.real ** 2 + psi.imag ** 2
    idx = torch.arange(psi.shape[1], device=psi.device)
    signs = torch.stack([1.0 - 2.0 * ((idx >> q) & 1).to(prob.dtype) for q in range(n)], dim=1)
    return prob @ signs


def run_ansatz(psi, theta_per_state, entangler="cnot"):
    """theta_per_state: [B, L, n, 2]."""
    _, L, n, _ = theta_per_state.shape
    for l in range(L):
        for q in range(n):
            psi = apply_rot(psi, "rx", theta_per_state[:, l, q, 0], q)
            psi = apply_rot(psi, "rz", theta_per_state[:, l, q, 1], q)
        for (c, t) in ring_pairs(n):
            psi = apply_cnot(psi, c, t) if entangler == "cnot" else apply_cz(psi, c, t)
    return psi


def vqc_state_ref(angles, theta, gid=None, basis="ry", entangler="cnot", init_state=None):
    """angles [B,n]; theta [G,L,n,2]; gid [B] int64 (None -> group 0)."""
    batch = angles.shape[0] if angles is not None else init_state.shape[0]
    th = theta[gid] if gid is not None else theta[:1].expand(batch, *theta.shape[1:])
    psi = product_state(angles, basis) if init_state is None else init_state.to(_cdtype(theta.dtype))
    if psi.real.dtype != th.dtype:
        th = th.to(psi.real.dtype)
    return run_ansatz(psi, th, entangler)


def vqc_expvals_ref(angles, theta, gid=None, basis="ry", entangler="cnot", init_state=None):
    return expval_z_all(vqc_state_ref(angles, theta, gid, basis, entangler, init_state))


def ansatz_unitary_ref(theta_one: torch.Tensor, entangler="cnot") -> torch.Tensor:
    """[L,n,2] -> dense [2^n, 2^n] unitary of the ansatz (columns = images of basis states)."""
    L, n, _ = theta_one.shape
    dim = 1 << n
    eye = torch.eye(dim, dtype=_cdtype(theta_one.dtype), device=theta_one.device)
    out = run_ansatz(eye, theta_one[None].expand(dim, L, n, 2), entangler)  # row b = U|b>
    return out.transpose(0, 1).contiguous()
