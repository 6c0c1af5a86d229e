"""Sharded-statevector planner: pure rank/qubit-map logic, checked on the CPU for every world size by
executing the plan on a dense NumPy state (physical qubit order) against the direct simulation."""
import numpy as np
import pytest

from qfedx_b200.ops import _lib
from qfedx_b200.ops.large import GATE, TWO_QUBIT, ansatz_gates
from qfedx_b200.oracle import statevector as sv
from qfedx_b200.parallel.sharded import plan_sharded

pytestmark = pytest.mark.skipif(not _lib.available(), reason="CUDA library not built")


def _apply(psi, kind, q0, q1, p, theta, angles):
    if kind == GATE["rx"]:
        return sv.apply_1q(psi, sv.gate_matrix("rx", theta[p]), q0)
    if kind == GATE["rz"]:
        return sv.apply_1q(psi, sv.gate_matrix("rz", theta[p]), q0)
    if kind == GATE["ry_feature"]:
        return sv.apply_1q(psi, sv.gate_matrix("ry", angles[p]), q0)
    if kind == GATE["cnot"]:
        return sv.apply_cnot(psi, q0, q1)
    if kind == GATE["cz"]:
        return sv.apply_cz(psi, q0, q1)
    raise AssertionError(kind)


def _swap_bits(psi, a, b):
    idx = np.arange(psi.size)
    ba, bb = (idx >> a) & 1, (idx >> b) & 1
    src = idx ^ ((ba ^ bb) << a) ^ ((ba ^ bb) << b)
    return psi[src]


@pytest.mark.parametrize("g_bits", [1, 2, 3])
@pytest.mark.parametrize("ent", ["cnot", "cz"])
def test_plan_execution_equals_direct_simulation(g_bits, ent):
    n, L = 9, 2
    rng = np.random.default_rng(g_bits)
    theta, angles = rng.normal(size=2 * n * L), rng.uniform(0, np.pi, n)
    gates = ansatz_gates(n, L, ent)
    plan = plan_sharded(n, g_bits, gates, compile_local=False)
    n_local = n - g_bits
    psi = sv.zero_state(n)
    for kind, payload in plan.steps:
        if kind == "local":
            for k, q0, q1, p in payload:
                assert 0 <= q0 < n_local and (k not in TWO_QUBIT or 0 <= q1 < n_local)   # never touches a global qubit
                psi = _apply(psi, k, q0, q1, p, theta, angles)
        else:
            for i in range(g_bits):
                psi = _swap_bits(psi, n_local + i, n_local - g_bits + i)
            for k, b0, b1, p0, p1 in payload.tolist():
                assert 0 <= b0 < g_bits and (k not in TWO_QUBIT or 0 <= b1 < g_bits)
                psi = _apply(psi, k, n_local - g_bits + b0, n_local - g_bits + b1, p0, theta, angles)
    # back to logical order
    idx = np.arange(psi.size)
    src = np.zeros_like(idx)
    for q, p in enumerate(plan.final_phys_of):
        src |= ((idx >> q) & 1) << p
    direct = sv.run_ansatz(sv.product_state_from_angles(angles), theta.reshape(L, n, 2), ent)
    np.testing.assert_allclose(psi[src], direct, atol=1e-12)
    assert sorted(plan.final_phys_of) == list(range(n))
    assert 1 <= plan.n_swaps <= 3 * L + 1


def test_fused_gates_follow_the_swap():
    # rotations on the global qubits are absorbed into the swap that brings them in
    n, g = 12, 2
    gates = [(GATE["rx"], q, -1, q) for q in range(n)]
    plan = plan_sharded(n, g, gates, compile_local=False)
    kinds = [k for k, _ in plan.steps]
    assert kinds == ["local", "swap"]
    assert len(plan.steps[1][1]) == g


def test_30_qubit_ring_plan_is_two_swaps_per_layer():
    plan = plan_sharded(30, 3, ansatz_gates(30, 3), compile_local=False)
    assert plan.n_swaps <= 2 * 3 + 1


def _run_plan(plan, n, g_bits, theta, angles):
    n_local = n - g_bits
    psi = sv.zero_state(n)
    for q in range(n):                                               # generic start state
        psi = sv.apply_1q(psi, sv.gate_matrix("ry", 0.4 + 0.15 * q), q)
    start = psi.copy()
    for kind, payload in plan.steps:
        if kind == "local":
            for k, q0, q1, p in payload:
                assert 0 <= q0 < n_local and (k not in TWO_QUBIT or 0 <= q1 < n_local)
                psi = _apply(psi, k, q0, q1, p, theta, angles)
        else:
            for i in range(g_bits):
                psi = _swap_bits(psi, n_local + i, n_local - g_bits + i)
            for k, b0, b1, p0, p1 in payload.tolist():
                psi = _apply(psi, k, n_local - g_bits + b0, n_local - g_bits + b1, p0, theta, angles)
    idx = np.arange(psi.size)
    src = np.zeros_like(idx)
    for q, p in enumerate(plan.final_phys_of):
        src |= ((idx >> q) & 1) << p
    return start, psi[src]


@pytest.mark.timeout(120)
@pytest.mark.parametrize("seed", range(10))
def test_random_circuits_through_the_sharded_planner(seed):
    """Arbitrary gate lists with long-range two-qubit gates (also between two global qubits, or a global and a
    top-local one - the case that once looped forever): the plan terminates and reproduces the circuit."""
    rng = np.random.default_rng(500 + seed)
    n, g_bits = int(rng.integers(7, 11)), int(rng.integers(1, 4))
    gates, n_params = [], 0
    for _ in range(int(rng.integers(10, 60))):
        kind = rng.choice(["rx", "rz", "cnot", "cz"])
        q = int(rng.integers(0, n))
        if kind in ("cnot", "cz"):
            t = int(rng.integers(0, n - 1))
            t += t >= q
            gates.append((GATE[kind], q, t, -1))
        else:
            gates.append((GATE[kind], q, -1, n_params))
            n_params += 1
    theta = rng.normal(size=max(1, n_params))
    plan = plan_sharded(n, g_bits, gates, compile_local=False)
    # note: the physical start state must be the logical one - the initial layout is the identity
    start, got = _run_plan(plan, n, g_bits, theta, None)
    ref = start
    for k, q0, q1, p in gates:
        ref = _apply(ref, k, q0, q1, p, theta, None)
    np.testing.assert_allclose(got, ref, atol=1e-12)
