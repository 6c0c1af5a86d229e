"""NumPy complex128 statevector oracle.

This is the correctness oracle for every statevector kernel in the framework and
the timed CPU baseline for the 4-qubit config.  qiskit cannot be installed in the
build image, so the gate conventions below restate qiskit's documented semantics
(SURVEY.md section 2.6): little-endian qubit order (qubit 0 is the least
significant bit of the basis index) and

    RX(a) = [[cos a/2, -i sin a/2], [-i sin a/2, cos a/2]]
    RY(a) = [[cos a/2,   -sin a/2], [   sin a/2, cos a/2]]
    RZ(a) = diag(exp(-i a/2), exp(+i a/2))

Reference call sites this stands in for: ``Statevector.from_instruction``
(/root/reference/src/QFed/qAmplitude.py:44-46) and the ``qc.rx/ry/rz`` layer
(/root/reference/src/QFed/qAngle.py:43-50).  The ansatz follows
/root/reference/ROADMAP.md:126-128.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "zero_state", "gate_matrix", "apply_1q", "apply_cnot", "apply_cz",
    "expval_z", "product_state_from_angles", "run_ansatz", "vqc_expvals",
    "ring_pairs", "GENERATORS",
]

_I = np.eye(2, dtype=np.complex128)
_X = np.array([[0, 1], [1, 0]], dtype=np.complex128)
_Y = np.array([[0, -1j], [1j, 0]], dtype=np.complex128)
_Z = np.array([[1, 0], [0, -1]], dtype=np.complex128)
GENERATORS = {"rx": _X, "ry": _Y, "rz": _Z}


def zero_state(n_qubits: int) -> np.ndarray:
    psi = np.zeros(1 << n_qubits, dtype=np.complex128)
    psi[0] = 1.0
    return psi


def gate_matrix(kind: str, angle: float) -> np.ndarray:
    """2x2 matrix of exp(-i angle P / 2) for P in {X, Y, Z}."""
    c, s = np.cos(angle / 2.0), np.sin(angle / 2.0)
    return c * _I - 1j * s * GENERATORS[kind.lower()]


def apply_1q(psi: np.ndarray, mat: np.ndarray, q: int) -> np.ndarray:
    """Apply a 2x2 matrix to qubit ``q`` (little-endian) of ``psi``; returns a new array."""
    n = int(np.log2(psi.size))
    # reshape so that axis 1 is the target bit: index = hi * 2^(q+1) + bit * 2^q + lo
    v = psi.reshape(1 << (n - q - 1), 2, 1 << q)
    out = np.einsum("ab,hbl->hal", mat, v)
    return out.reshape(-1)


def apply_cnot(psi: np.ndarray, control: int, target: int) -> np.ndarray:
    idx = np.arange(psi.size)
    src = np.where((idx >> control) & 1, idx ^ (1 << target), idx)
    return psi[src]


def apply_cz(psi: np.ndarray, a: int, b: int) -> np.ndarray:
    idx = np.arange(psi.size)
    sign = np.where(((idx >> a) & 1) & ((idx >> b) & 1), -1.0, 1.0)
    return psi * sign


def expval_z(psi: np.ndarray, q: int) -> float:
    idx = np.arange(psi.size)
    sign = 1.0 - 2.0 * ((idx >> q) & 1)
    return float(np.sum(sign * (psi.real ** 2 + psi.imag ** 2)))


def product_state_from_angles(angles: np.ndarray, basis: str = "ry") -> np.ndarray:
    """State after one rotation per qubit on |0..0> (the angle feature map)."""
    psi = zero_state(len(angles))
    for q, a in enumerate(angles):
        psi = apply_1q(psi, gate_matrix(basis, float(a)), q)
    return psi


def ring_pairs(n_qubits: int):
    """Entangler pairs (control, target) applied in order after the rotations."""
    if n_qubits < 2:
        return []
    if n_qubits == 2:
        return [(0, 1)]
    return [(q, (q + 1) % n_qubits) for q in range(n_qubits)]


def run_ansatz(psi: np.ndarray, theta: np.ndarray, entangler: str = "cnot") -> np.ndarray:
    """Hardware-efficient ansatz: per layer RX(theta[l,q,0]) RZ(theta[l,q,1]) on every
    qubit, then the ring entangler.  ``theta`` has shape [L, n, 2]."""
    n_layers, n, _ = theta.shape
    for l in range(n_layers):
        for q in range(n):
            psi = apply_1q(psi, gate_matrix("rx", theta[l, q, 0]), q)
            psi = apply_1q(psi, gate_matrix("rz", theta[l, q, 1]), q)
        for (c, t) in ring_pairs(n):
            psi = apply_cnot(psi, c, t) if entangler == "cnot" else apply_cz(psi, c, t)
    return psi


def vqc_expvals(angles: np.ndarray, theta: np.ndarray, basis: str = "ry",
                entangler: str = "cnot", init_state: np.ndarray | None = None) -> np.ndarray:
    """<Z_q> for every qubit after feature map + ansatz.  ``angles`` [n] are the
    already-scaled feature-map rotation angles; ``init_state`` overrides the feature
    map with an amplitude-encoded state."""
    n = theta.shape[1]
    psi = product_state_from_angles(angles, basis) if init_state is None else np.asarray(
        init_state, dtype=np.complex128)
    psi = run_ansatz(psi, np.asarray(theta, dtype=np.float64), entangler)
    return np.array([expval_z(psi, q) for q in range(n)])
