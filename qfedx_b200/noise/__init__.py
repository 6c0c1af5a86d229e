"""Noise models for VQC simulation (/root/reference/ROADMAP.md:64-73: depolarizing, amplitude damping,
readout error, finite shots).  The statevector engine is pure-state, so channels are applied where
they have an exact closed form on the measured quantities, and by Monte-Carlo trajectories otherwise:

* depolarizing noise on the readout qubit after the circuit, readout confusion and finite shots act
  on <Z> exactly (``apply_to_expvals``): <Z> -> (1-p)<Z>, then the confusion matrix, then a binomial
  sample of ``shots`` measurements.
* per-gate depolarizing / amplitude damping inside the circuit: ``trajectory_expvals`` samples Kraus
  branches on the reference simulator (ops/ref.py) and averages <Z> over trajectories.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from ..ops import ref
from ..oracle.statevector import ring_pairs


@dataclass
class NoiseConfig:
    kind: str = "none"             # none | depolarizing | amplitude
    p: float = 0.0                 # depolarizing probability per gate / on the readout
    gamma: float = 0.0             # amplitude-damping probability per gate
    readout_p01: float = 0.0       # P(read 1 | state 0)
    readout_p10: float = 0.0       # P(read 0 | state 1)
    shots: int = 0                 # 0 = exact expectation values

    @property
    def enabled(self) -> bool:
        return self.kind != "none" or self.shots > 0 or self.readout_p01 > 0 or self.readout_p10 > 0


def readout_affine(cfg: NoiseConfig):
    """(alpha, beta) with E' = alpha * E + beta for the closed-form channels acting on a measured <Z>: depolarizing
    on the readout qubit, amplitude damping towards |0>, then the readout confusion matrix.  The sm_100a training and
    evaluation kernels apply exactly this map in their <Z> epilogue (csrc/common.cuh: ReadoutNoise)."""
    a, b = 1.0, 0.0
    if cfg.kind == "depolarizing" and cfg.p > 0:
        a, b = (1.0 - cfg.p) * a, (1.0 - cfg.p) * b
    if cfg.kind == "amplitude" and cfg.gamma > 0:
        a, b = (1.0 - cfg.gamma) * a, (1.0 - cfg.gamma) * b + cfg.gamma
    if cfg.readout_p01 > 0 or cfg.readout_p10 > 0:
        s, t = 1.0 - cfg.readout_p01 - cfg.readout_p10, cfg.readout_p10 - cfg.readout_p01
        a, b = s * a, s * b + t
    return a, b


def noisy_expvals_cuda(E: torch.Tensor, cfg: NoiseConfig, key: int = 0) -> torch.Tensor:
    """Readout noise on CUDA expectation values through the library's K3 epilogue pass (no autograd)."""
    from ..ops import _lib
    a, b = readout_affine(cfg)
    E = E.contiguous().float()
    out = torch.empty_like(E)
    _lib.call("qfx_readout_noise", _lib.ptr(E), _lib.ptr(out), _lib.c_int(E.shape[0]), _lib.c_int(E.shape[1]),
              _lib.c_float(a), _lib.c_float(b), _lib.c_int(int(cfg.shots)), _lib.c_u64(key & (2 ** 64 - 1)),
              _lib.stream_ptr())
    return out


def apply_to_expvals(E: torch.Tensor, cfg: NoiseConfig, generator=None) -> torch.Tensor:
    """Readout-level noise on exact <Z> values [B, n] (differentiable except for shot sampling)."""
    out = E
    if cfg.kind == "depolarizing" and cfg.p > 0:
        out = (1.0 - cfg.p) * out
    if cfg.kind == "amplitude" and cfg.gamma > 0:
        out = (1.0 - cfg.gamma) * out + cfg.gamma           # decay towards |0>: <Z> -> (1-g)<Z> + g
    if cfg.readout_p01 > 0 or cfg.readout_p10 > 0:
        p0 = 0.5 * (1.0 + out)                               # P(state 0)
        read0 = p0 * (1.0 - cfg.readout_p01) + (1.0 - p0) * cfg.readout_p10
        out = 2.0 * read0 - 1.0
    if cfg.shots > 0:
        p0 = (0.5 * (1.0 + out.detach())).clamp(0, 1)
        k = torch.binomial(torch.full_like(p0, float(cfg.shots)), p0, generator=generator)
        sampled = 2.0 * k / cfg.shots - 1.0
        out = out + (sampled - out).detach()                 # straight-through: value sampled, gradient exact
    return out


def _kraus_step(psi, q, cfg: NoiseConfig, gen):
    """One stochastic Kraus branch on qubit q for every state of the batch."""
    B = psi.shape[0]
    if cfg.kind == "depolarizing" and cfg.p > 0:
        u = torch.rand(B, generator=gen, device=psi.device)
        which = torch.zeros(B, dtype=torch.long, device=psi.device)
        which[u < cfg.p] = torch.randint(1, 4, (int((u < cfg.p).sum()),), generator=gen, device=psi.device)
        half_pi = torch.full((B,), torch.pi, dtype=psi.real.dtype, device=psi.device)
        for pauli, kind in ((1, "rx"), (2, "ry"), (3, "rz")):   # a pi rotation is the Pauli up to a phase
            sel = which == pauli
            if sel.any():
                psi = torch.where(sel[:, None], ref.apply_rot(psi, kind, half_pi, q), psi)
    elif cfg.kind == "amplitude" and cfg.gamma > 0:
        dim = psi.shape[1]
        idx = torch.arange(dim, device=psi.device)
        one = ((idx >> q) & 1).bool()
        p1 = (psi.abs() ** 2)[:, one].sum(1)
        jump = torch.rand(B, generator=gen, device=psi.device) < cfg.gamma * p1
        damp = torch.where(one, torch.sqrt(torch.tensor(1.0 - cfg.gamma)), torch.tensor(1.0)).to(psi.real.dtype)
        no_jump = psi * damp
        jumped = torch.zeros_like(psi)
        jumped[:, ~one] = psi[:, one]                         # |1> -> |0>
        psi = torch.where(jump[:, None], jumped, no_jump)
        psi = psi / psi.norm(dim=1, keepdim=True).clamp_min(1e-20)
    return psi


def trajectory_expvals(angles, theta, cfg: NoiseConfig, n_traj: int = 64, entangler="cnot", generator=None):
    """Average <Z_q> over stochastic trajectories with a noise event after every rotation pair."""
    th = theta if theta.dim() == 4 else theta[None]
    L, n = th.shape[1], th.shape[2]
    acc = 0.0
    for _ in range(n_traj):
        psi = ref.product_state(angles)
        tb = th[:1].expand(angles.shape[0], L, n, 2)
        for l in range(L):
            for q in range(n):
                psi = ref.apply_rot(psi, "rx", tb[:, l, q, 0], q)
                psi = ref.apply_rot(psi, "rz", tb[:, l, q, 1], q)
                psi = _kraus_step(psi, q, cfg, generator)
            for c, t in ring_pairs(n):
                psi = ref.apply_cnot(psi, c, t) if entangler == "cnot" else ref.apply_cz(psi, c, t)
        acc = acc + ref.expval_z_all(psi)
    return apply_to_expvals(acc / n_traj, NoiseConfig(readout_p01=cfg.readout_p01, readout_p10=cfg.readout_p10,
                                                      shots=cfg.shots), generator)
