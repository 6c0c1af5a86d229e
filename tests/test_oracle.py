"""Pins the CPU oracle to hand-computed cases (qiskit is not installable) and checks
the three gradient methods against each other (ROADMAP check: parameter-shift == adjoint)."""
import math

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from qfedx_b200.oracle import encoders as enc
from qfedx_b200.oracle import statevector as sv
from qfedx_b200.oracle import vqc as ov


def test_single_qubit_rotations_from_zero():
    a = 0.7
    c, s = math.cos(a / 2), math.sin(a / 2)
    np.testing.assert_allclose(sv.apply_1q(sv.zero_state(1), sv.gate_matrix("ry", a), 0), [c, s], atol=1e-15)
    np.testing.assert_allclose(sv.apply_1q(sv.zero_state(1), sv.gate_matrix("rx", a), 0), [c, -1j * s], atol=1e-15)
    np.testing.assert_allclose(sv.apply_1q(sv.zero_state(1), sv.gate_matrix("rz", a), 0),
                               [np.exp(-1j * a / 2), 0], atol=1e-15)


def test_little_endian_product_state():
    # qubit 0 = pi (|1>), qubit 1 = 0 (|0>)  ->  |q1 q0> = |01> = index 1
    psi = sv.product_state_from_angles(np.array([math.pi, 0.0]))
    np.testing.assert_allclose(np.abs(psi), [0, 1, 0, 0], atol=1e-15)
    assert sv.expval_z(psi, 0) == pytest.approx(-1.0)
    assert sv.expval_z(psi, 1) == pytest.approx(1.0)


def test_cnot_makes_bell_state():
    psi = sv.apply_1q(sv.zero_state(2), sv.gate_matrix("ry", math.pi / 2), 0)  # (|0>+|1>)/sqrt2 on q0
    psi = sv.apply_cnot(psi, 0, 1)
    np.testing.assert_allclose(psi, np.array([1, 0, 0, 1]) / math.sqrt(2), atol=1e-15)
    assert sv.expval_z(psi, 0) == pytest.approx(0.0, abs=1e-15)


def test_cz_sign():
    psi = np.ones(4, dtype=np.complex128) / 2
    np.testing.assert_allclose(sv.apply_cz(psi, 0, 1), np.array([1, 1, 1, -1]) / 2)


def test_ry_expectation_closed_form():
    # <Z> after RY(a) is cos(a)
    for a in (0.0, 0.3, 1.2, math.pi):
        psi = sv.product_state_from_angles(np.array([a]))
        assert sv.expval_z(psi, 0) == pytest.approx(math.cos(a), abs=1e-14)


def test_ring_pairs():
    assert sv.ring_pairs(1) == []
    assert sv.ring_pairs(2) == [(0, 1)]
    assert sv.ring_pairs(4) == [(0, 1), (1, 2), (2, 3), (3, 0)]


@settings(max_examples=15, deadline=None)
@given(n=st.integers(1, 5), layers=st.integers(1, 3), seed=st.integers(0, 10 ** 6),
       ent=st.sampled_from(["cnot", "cz"]))
def test_gradients_agree(n, layers, seed, ent):
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, math.pi, n)
    theta = rng.normal(size=(layers, n, 2))
    cot = rng.normal(size=n)
    ga, gang = ov.grad_adjoint(ang, theta, cot, entangler=ent, want_angle_grad=True)
    gs = ov.grad_param_shift(ang, theta, cot, entangler=ent)
    gf = ov.grad_finite_diff(ang, theta, cot, entangler=ent)
    np.testing.assert_allclose(ga, gs, atol=1e-10)
    np.testing.assert_allclose(ga, gf, atol=1e-6)
    # feature-map angle gradient by finite differences
    eps = 1e-6
    for q in range(n):
        ap, am = ang.copy(), ang.copy()
        ap[q] += eps; am[q] -= eps
        fd = float(np.dot(cot, sv.vqc_expvals(ap, theta, entangler=ent) - sv.vqc_expvals(am, theta, entangler=ent))) / (2 * eps)
        assert gang[q] == pytest.approx(fd, abs=1e-6)


def test_state_stays_normalised():
    rng = np.random.default_rng(0)
    psi = sv.run_ansatz(sv.product_state_from_angles(rng.uniform(0, 3, 6)), rng.normal(size=(4, 6, 2)))
    assert np.vdot(psi, psi).real == pytest.approx(1.0, abs=1e-12)


def test_amplitude_state_and_uniform_fallback():
    v = np.arange(16, dtype=float)
    st_ = enc.amplitude_state(v)
    np.testing.assert_allclose(st_, v / np.linalg.norm(v))
    np.testing.assert_allclose(enc.normalize_for_amplitude(np.zeros(8)), np.full(8, 1 / math.sqrt(8)))
    with pytest.raises(ValueError):
        enc.amplitude_state(np.ones(6))


def test_angle_features_minmax_and_pool():
    f = enc.angle_features(np.array([2.0, 4.0, 3.0, 2.0]))
    np.testing.assert_allclose(f, [0, math.pi, math.pi / 2, 0])
    np.testing.assert_allclose(enc.angle_features(np.ones(4)), np.zeros(4))
    np.testing.assert_allclose(enc.pool_to_n_features(np.arange(10.0), 4), [0.5, 2.5, 4.5, 7.5])
    np.testing.assert_allclose(enc.pool_to_n_features(np.array([1.0, 2.0]), 4), [1, 2, 0, 0])


def test_downsample_is_7x7_block_mean():
    img = np.arange(28 * 28, dtype=float).reshape(28, 28)
    out = enc.downsample_image(img)
    assert out.shape == (4, 4)
    assert out[1, 2] == pytest.approx(img[7:14, 14:21].mean())


def test_loss_gradient_matches_finite_difference():
    rng = np.random.default_rng(3)
    n, L = 3, 2
    ang, theta = rng.uniform(0, 3, n), rng.normal(size=(L, n, 2))
    a, b = np.array([1.3]), np.array([-0.2])
    E = sv.vqc_expvals(ang, theta)
    cot = ov.dloss_dexp(E, a, b, [0], 1, n)
    g = ov.grad_adjoint(ang, theta, cot)
    eps = 1e-6
    tp, tm = theta.copy(), theta.copy()
    tp[1, 2, 0] += eps; tm[1, 2, 0] -= eps
    fd = (ov.vqc_loss(ang, tp, a, b, [0], 1) - ov.vqc_loss(ang, tm, a, b, [0], 1)) / (2 * eps)
    assert g[1, 2, 0] == pytest.approx(fd, abs=1e-6)
