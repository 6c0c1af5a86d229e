"""Oracle VQC: loss and three independent gradient methods in float64.

* ``grad_param_shift`` - two-term shift rule d<O>/dt = 0.5(<O>(t+pi/2) - <O>(t-pi/2))
  (/root/reference/ROADMAP.md:130-135).
* ``grad_adjoint``     - reverse sweep carrying |psi> and |lambda>.
* ``grad_finite_diff`` - central differences, the tie-breaker.

The planned check "gradients from parameter-shift match adjoint within tolerance"
(/root/reference/ROADMAP.md:27) is tests/test_oracle.py.
"""
from __future__ import annotations

import numpy as np

from . import statevector as sv

__all__ = ["logits_from_expvals", "loss_from_logits", "vqc_loss", "grad_param_shift",
           "grad_adjoint", "grad_finite_diff", "dloss_dexp"]


def logits_from_expvals(E: np.ndarray, a: np.ndarray, b: np.ndarray, readout) -> np.ndarray:
    return a * E[..., list(readout)] + b


def loss_from_logits(logits: np.ndarray, y: int) -> float:
    """Binary (one logit): BCE-with-logits on label y in {0,1}; multiclass: softmax CE."""
    if logits.shape[-1] == 1:
        z = float(logits[0])
        return float(max(z, 0.0) - z * y + np.log1p(np.exp(-abs(z))))
    m = logits.max()
    lse = m + np.log(np.exp(logits - m).sum())
    return float(lse - logits[y])


def dloss_dlogits(logits: np.ndarray, y: int) -> np.ndarray:
    if logits.shape[-1] == 1:
        return np.array([1.0 / (1.0 + np.exp(-logits[0])) - y])
    p = np.exp(logits - logits.max())
    p /= p.sum()
    p[y] -= 1.0
    return p


def dloss_dexp(E, a, b, readout, y, n):
    """Cotangent of the loss w.r.t. <Z_q> for every qubit (zero off the readout set)."""
    g = np.zeros(n)
    dl = dloss_dlogits(logits_from_expvals(E, a, b, readout), y)
    for j, q in enumerate(readout):
        g[q] += dl[j] * a[j]
    return g


def vqc_loss(angles, theta, a, b, readout, y, basis="ry", entangler="cnot") -> float:
    E = sv.vqc_expvals(angles, theta, basis, entangler)
    return loss_from_logits(logits_from_expvals(E, a, b, readout), y)


def grad_param_shift(angles, theta, cot, basis="ry", entangler="cnot") -> np.ndarray:
    """d(sum_q cot[q] <Z_q>)/d theta by the two-term shift rule."""
    theta = np.asarray(theta, dtype=np.float64)
    g = np.zeros_like(theta)
    for idx in np.ndindex(theta.shape):
        tp, tm = theta.copy(), theta.copy()
        tp[idx] += np.pi / 2
        tm[idx] -= np.pi / 2
        ep = sv.vqc_expvals(angles, tp, basis, entangler)
        em = sv.vqc_expvals(angles, tm, basis, entangler)
        g[idx] = 0.5 * float(np.dot(cot, ep - em))
    return g


def grad_finite_diff(angles, theta, cot, basis="ry", entangler="cnot", eps=1e-6) -> np.ndarray:
    theta = np.asarray(theta, dtype=np.float64)
    g = np.zeros_like(theta)
    for idx in np.ndindex(theta.shape):
        tp, tm = theta.copy(), theta.copy()
        tp[idx] += eps
        tm[idx] -= eps
        ep = sv.vqc_expvals(angles, tp, basis, entangler)
        em = sv.vqc_expvals(angles, tm, basis, entangler)
        g[idx] = float(np.dot(cot, ep - em)) / (2 * eps)
    return g


def grad_adjoint(angles, theta, cot, basis="ry", entangler="cnot", want_angle_grad=False):
    """Adjoint differentiation of sum_q cot[q] <Z_q>.

    With psi_j the state after gate j and lambda_j = (U_{>j})^dagger O psi_final,
    d<O>/d theta_j = Im <lambda_j| P_j |psi_j> for a gate exp(-i theta_j P_j / 2).
    """
    theta = np.asarray(theta, dtype=np.float64)
    n_layers, n, _ = theta.shape
    psi = sv.run_ansatz(sv.product_state_from_angles(angles, basis), theta, entangler)
    idx = np.arange(psi.size)
    diag = np.zeros(psi.size)
    for q in range(n):
        diag += cot[q] * (1.0 - 2.0 * ((idx >> q) & 1))
    lam = diag * psi
    g = np.zeros_like(theta)

    def inner_grad(P, q):
        return float(np.imag(np.vdot(lam, sv.apply_1q(psi, P, q))))

    for l in reversed(range(n_layers)):
        for (c, t) in reversed(sv.ring_pairs(n)):
            if entangler == "cnot":
                psi, lam = sv.apply_cnot(psi, c, t), sv.apply_cnot(lam, c, t)
            else:
                psi, lam = sv.apply_cz(psi, c, t), sv.apply_cz(lam, c, t)
        for q in reversed(range(n)):
            for k, kind in ((1, "rz"), (0, "rx")):
                g[l, q, k] = inner_grad(sv.GENERATORS[kind], q)
                inv = sv.gate_matrix(kind, -theta[l, q, k])
                psi, lam = sv.apply_1q(psi, inv, q), sv.apply_1q(lam, inv, q)
    if not want_angle_grad:
        return g
    ga = np.zeros(n)
    for q in range(n):
        ga[q] = inner_grad(sv.GENERATORS[basis], q)
    return g, ga
