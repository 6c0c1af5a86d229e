"""Pass compiler (host C++) - runs without a GPU."""
import numpy as np
import pytest
import torch

from qfedx_b200.ops import _lib

pytestmark = pytest.mark.skipif(not _lib.available(), reason="CUDA library not built")


def _check_schedule(n, gates, passes, tile_bits, low_bits):
    from qfedx_b200.ops.large import GATE, TWO_QUBIT
    # every pass: ascending tile qubits, starts with the low qubits, ops only use tile bits
    seen = []
    for p in passes:
        tq = p.tile_qubits.tolist()
        assert tq == sorted(tq) and len(set(tq)) == len(tq) == min(tile_bits, n)
        for kind, b0, b1, p0, p1 in p.ops.tolist():
            assert 0 <= b0 < len(tq)
            q0 = tq[b0]
            q1 = tq[b1] if kind in TWO_QUBIT else -1
            seen.append((kind, q0, q1, p0, p1))
    # per-qubit order of the original program is preserved (after RX+RZ fusion)
    def per_qubit(seq):
        out = {}
        for kind, q0, q1, p0, p1 in seq:
            params = [p for p in (p0, p1) if p >= 0] or [-1]
            for q in (q0, q1):
                if q >= 0:
                    out.setdefault(q, []).extend((kind if kind != GATE["rxrz"] else "f", p) for p in params)
        return out
    orig = []
    for kind, q0, q1, p in gates:
        orig.append((kind, q0, q1, p, -1))
    a, b = per_qubit(orig), per_qubit(seen)
    for q in a:
        pa = [p for _, p in a[q]]
        pb = [p for _, p in b[q]]
        assert pa == pb, (q, pa, pb)


def test_ring_ansatz_20q_two_passes_per_layer():
    from qfedx_b200.ops.large import ansatz_gates, compile_passes
    gates = ansatz_gates(20, 3)
    passes = compile_passes(20, gates, 12, 3)
    _check_schedule(20, gates, passes, 12, 3)
    assert len(passes) <= 2 * 3 + 2
    n_ops = sum(len(p.ops) for p in passes)
    assert n_ops == 20 + 3 * (20 + 20)        # feature + per layer (20 fused RX.RZ + 20 CNOT)


@pytest.mark.parametrize("n,L,tile,low", [(5, 2, 4, 2), (5, 2, 4, 3), (13, 2, 12, 3), (16, 1, 10, 3), (3, 1, 12, 3)])
def test_schedules_are_valid(n, L, tile, low):
    from qfedx_b200.ops.large import ansatz_gates, compile_passes
    gates = ansatz_gates(n, L, "cz")
    passes = compile_passes(n, gates, tile, low)
    _check_schedule(n, gates, passes, min(tile, n), min(low, n))


def test_bad_gate_is_rejected():
    from qfedx_b200.ops.large import compile_passes
    with pytest.raises(_lib.QfxError):
        compile_passes(4, [(0, 7, -1, 0)], 4, 2)


def test_segment_plan_is_a_valid_reordering():
    """Every op appears once, only on active positions, per-bit order preserved; one-qubit gates and CNOT targets
    sit on register positions; four of the five lane bits cover the classes mod 4 (conflict-free re-layout)."""
    from qfedx_b200.ops.large import TWO_QUBIT, ansatz_gates, compile_passes
    passes = compile_passes(20, ansatz_gates(20, 3), 12, 3)
    for p in passes:
        plan = p.plan.tolist()
        assert plan[0] >> 16 == 1                                                      # trailer present
        n_seg, cur, seen = plan[0] & 0xFFFF, 1, []
        for _ in range(n_seg):
            p2b, n_ops = plan[cur:cur + 12], plan[cur + 12]
            assert sorted(p2b) == list(range(12)) and sorted(b % 4 for b in p2b[:4]) == [0, 1, 2, 3]
            for o in range(n_ops):
                kind, pa, pb, idx = plan[cur + 13 + 4 * o: cur + 17 + 4 * o]
                assert 0 <= pa < 9 and (kind not in TWO_QUBIT or 0 <= pb < 9)          # never a warp (inactive) bit
                assert p2b[pa] == p.ops[idx][1] and (kind not in TWO_QUBIT or p2b[pb] == p.ops[idx][2])
                if kind == 3:
                    assert 5 <= pb < 9                                                  # CNOT target in registers
                elif kind != 4:
                    assert 5 <= pa < 9                                                  # one-qubit gate in registers
                seen.append(idx)
            cur += 13 + 4 * n_ops
        assert sorted(seen) == list(range(len(p.ops)))
        # trailer: op modes + carrier; phase-stripped gates need a carrier to absorb their scalar phases
        modes, carrier = plan[cur:cur + len(p.ops)], plan[cur + len(p.ops)]
        assert cur + len(p.ops) + 1 == len(plan)
        assert all(p.ops[i][0] == 5 for i, m in enumerate(modes) if m)                  # only fused RX.RZ gates
        assert (carrier >= 0 and modes[carrier] == 2 and modes.count(2) == 1) if any(modes) else carrier == -1
        # per-bit order of the ops is unchanged
        for b in range(12):
            touching = [i for i in range(len(p.ops)) if b in (p.ops[i][1], p.ops[i][2] if p.ops[i][0] in TWO_QUBIT else -9)]
            assert [i for i in seen if i in touching] == touching
        assert n_seg <= 6


@pytest.mark.parametrize("n,ent", [(3, "cnot"), (6, "cnot"), (5, "cz")])
def test_ring_readout_masks_replace_the_final_ring(n, ent):
    """<Z_q> behind the last entangling ring == parity readout in front of it."""
    from qfedx_b200.ops import ref
    from qfedx_b200.ops.large import ring_readout_masks
    from qfedx_b200.oracle.statevector import ring_pairs
    L = 2
    g = torch.Generator().manual_seed(n)
    ang = torch.rand(4, n, generator=g).double() * 3
    theta = torch.randn(1, L, n, 2, generator=g).double()
    full = ref.vqc_state_ref(ang, theta, None, "ry", ent)
    expect = ref.expval_z_all(full).numpy()
    idx = np.arange(1 << n)
    st = full.numpy()
    for c, t in reversed(ring_pairs(n)):          # undo the last ring (CNOT and CZ are involutions)
        if ent == "cnot":
            st = st[:, idx ^ (((idx >> c) & 1) << t)]
        else:
            st = st * np.where(((idx >> c) & 1) & ((idx >> t) & 1), -1.0, 1.0)
    p = np.abs(st) ** 2
    masks = ring_readout_masks(n, ent)
    got = np.stack([(p * np.where([bin(i & m).count("1") & 1 for i in idx], -1.0, 1.0)).sum(1) for m in masks], 1)
    np.testing.assert_allclose(got, expect, atol=1e-12)


def _emulate_pass_plan(psi, p, theta, angles):
    """NumPy restatement of what sv_large_pass_v2_kernel does with a segment plan (csrc/sv_large.cu): gates in plan
    order on the tile's qubits; phase-stripped RX.RZ gates apply diag(1, e^{i phi}) RX(theta) and the carrier gate
    absorbs the product of the dropped scalars e^{-i phi/2}."""
    from qfedx_b200.ops.large import GATE, PLAN_KIND_RXRZ_STRIPPED
    from qfedx_b200.oracle import statevector as sv
    plan, tq = p.plan.tolist(), [int(q) for q in p.tile_qubits]
    n_seg, cur = plan[0] & 0xFFFF, 1
    n_ops = len(p.ops)
    modes, carrier = plan[len(plan) - n_ops - 1:len(plan) - 1], plan[-1]
    dropped = sum(-0.5 * theta[int(p.ops[i][4])] for i in range(n_ops) if modes[i] == 1)
    for _ in range(n_seg):
        p2b, k_ops = plan[cur:cur + 12], plan[cur + 12]
        for o in range(k_ops):
            kind, pa, pb, idx = plan[cur + 13 + 4 * o: cur + 17 + 4 * o]
            op = p.ops[idx]
            q0 = tq[p2b[pa]]
            if kind == GATE["cnot"]:
                psi = sv.apply_cnot(psi, q0, tq[p2b[pb]])
            elif kind == GATE["cz"]:
                psi = sv.apply_cz(psi, q0, tq[p2b[pb]])
            elif kind == GATE["ry_feature"]:
                psi = sv.apply_1q(psi, sv.gate_matrix("ry", angles[int(op[3])]), q0)
            elif kind == PLAN_KIND_RXRZ_STRIPPED:
                assert modes[idx] == 1
                m = np.diag([1.0, np.exp(1j * theta[int(op[4])])]) @ sv.gate_matrix("rx", theta[int(op[3])])
                psi = sv.apply_1q(psi, m, q0)
            elif kind == GATE["rxrz"]:
                m = sv.gate_matrix("rz", theta[int(op[4])]) @ sv.gate_matrix("rx", theta[int(op[3])])
                if idx == carrier:
                    m = m * np.exp(1j * dropped)
                psi = sv.apply_1q(psi, m, q0)
            else:
                name = {GATE["rx"]: "rx", GATE["ry"]: "ry", GATE["rz"]: "rz"}[kind]
                psi = sv.apply_1q(psi, sv.gate_matrix(name, theta[int(op[3])]), q0)
        cur += 13 + 4 * k_ops
    return psi


@pytest.mark.parametrize("n,L,ent,final_ring", [(12, 2, "cnot", True), (13, 2, "cnot", False), (14, 3, "cz", True),
                                                (13, 1, "cnot", True)])
def test_segment_plans_reproduce_the_circuit(n, L, ent, final_ring):
    """Pass compiler + segment planner + phase folding, executed on a dense NumPy state, equal the direct simulation."""
    from qfedx_b200.ops.large import ansatz_gates, compile_passes
    from qfedx_b200.oracle import statevector as sv
    rng = np.random.default_rng(n * 10 + L)
    theta, angles = rng.normal(size=2 * n * L), rng.uniform(0, np.pi, n)
    passes = compile_passes(n, ansatz_gates(n, L, ent, final_ring=final_ring), 12, 3)
    assert all(p.plan is not None for p in passes)
    psi = sv.zero_state(n)
    for p in passes:
        psi = _emulate_pass_plan(psi, p, theta, angles)
    ref = sv.zero_state(n)
    for q in range(n):
        ref = sv.apply_1q(ref, sv.gate_matrix("ry", angles[q]), q)
    for l in range(L):
        for q in range(n):
            ref = sv.apply_1q(ref, sv.gate_matrix("rx", theta[(l * n + q) * 2]), q)
            ref = sv.apply_1q(ref, sv.gate_matrix("rz", theta[(l * n + q) * 2 + 1]), q)
        if ent != "none" and (final_ring or l < L - 1):
            for c, t in sv.ring_pairs(n):
                ref = sv.apply_cnot(ref, c, t) if ent == "cnot" else sv.apply_cz(ref, c, t)
    np.testing.assert_allclose(psi, ref, atol=1e-10)


def _random_circuit(rng, n, n_gates):
    from qfedx_b200.ops.large import GATE
    gates, n_params = [], 0
    for _ in range(n_gates):
        kind = rng.choice(["rx", "ry", "rz", "cnot", "cz", "rx+rz"])
        q = int(rng.integers(0, n))
        if kind in ("cnot", "cz"):
            t = int(rng.integers(0, n - 1))
            t += t >= q
            gates.append((GATE[kind], q, t, -1))
        elif kind == "rx+rz":                      # the pair the compiler fuses
            gates.append((GATE["rx"], q, -1, n_params))
            gates.append((GATE["rz"], q, -1, n_params + 1))
            n_params += 2
        else:
            gates.append((GATE[kind], q, -1, n_params))
            n_params += 1
    return gates, n_params


@pytest.mark.parametrize("seed", range(12))
def test_random_circuits_compile_and_plan_correctly(seed):
    """Arbitrary gate lists (long-range CNOT / CZ, unfused rotations): compiled passes + segment plans executed on a
    dense NumPy state equal the gate-by-gate simulation; every pass keeps the scheduling invariants."""
    from qfedx_b200.ops.large import GATE, compile_passes
    from qfedx_b200.oracle import statevector as sv
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(12, 15))
    gates, n_params = _random_circuit(rng, n, int(rng.integers(20, 90)))
    theta = rng.normal(size=max(1, n_params))
    passes = compile_passes(n, gates, 12, int(rng.integers(0, 4)))
    psi = sv.zero_state(n)
    psi = sv.apply_1q(psi, sv.gate_matrix("ry", 0.7), 0)          # not an eigenstate of anything
    for q in range(1, n):
        psi = sv.apply_1q(psi, sv.gate_matrix("rx", 0.3 + 0.1 * q), q)
    ref = psi.copy()
    for p in passes:
        assert p.plan is not None
        psi = _emulate_pass_plan(psi, p, theta, None)
    names = {GATE["rx"]: "rx", GATE["ry"]: "ry", GATE["rz"]: "rz"}
    for k, q0, q1, prm in gates:
        if k == GATE["cnot"]:
            ref = sv.apply_cnot(ref, q0, q1)
        elif k == GATE["cz"]:
            ref = sv.apply_cz(ref, q0, q1)
        else:
            ref = sv.apply_1q(ref, sv.gate_matrix(names[k], theta[prm]), q0)
    np.testing.assert_allclose(psi, ref, atol=1e-10)


def test_pass_counts_of_the_benchmark_circuits_do_not_regress():
    """Pass count = trips of the whole state through memory; these are the schedules the measured numbers rest on."""
    from qfedx_b200.ops.large import ansatz_gates, compile_passes
    from qfedx_b200.parallel.sharded import plan_sharded
    readout = lambda n, L: len(compile_passes(n, ansatz_gates(n, L, final_ring=False), 12, 3))
    assert readout(20, 2) == 2                      # config 4
    assert readout(24, 2) <= 3
    assert readout(27, 2) <= 3
    assert readout(30, 2) <= 4                      # config 5 on one GPU
    plan = plan_sharded(30, 3, ansatz_gates(30, 2, final_ring=False))
    assert plan.n_swaps == 1 and sum(len(p) for k, p in plan.steps if k == "local") <= 4   # config 5 on 8 GPUs


def test_lockstep_adjoint_training_equals_the_one_client_at_a_time_loop():
    """All clients in one batch of states (row s * K + k -> parameter row k) follow the per-client loop; coordinates
    whose gradient is zero up to rounding may differ by Adam-sized steps, everything else agrees."""
    import math
    from qfedx_b200.fl import FederatedDataset
    from qfedx_b200.fl.trainers_large import LargeVQCTrainer, PassBackend
    from qfedx_b200.models import VQC
    n, L = 5, 2
    g = torch.Generator().manual_seed(3)
    clients = [((torch.rand(7 + k, n, generator=g) * math.pi), torch.randint(0, 2, (7 + k,), generator=g)) for k in range(3)]
    data = FederatedDataset.from_clients(clients, "cpu")
    torch.manual_seed(0)
    model = VQC(n, L, 2)
    outs = []
    for mode in ("seq", "lock"):
        gen = torch.Generator().manual_seed(7)
        tr = LargeVQCTrainer(model, data, PassBackend(n, L), epochs=2, batch_size=3, lr=0.05, grad_method="adjoint", generator=gen)
        if mode == "seq":
            outs.append(torch.stack([tr._train_one(tr.global_flat(), k)[0] for k in range(3)]))
        else:
            outs.append(tr._train_lockstep(tr.global_flat(), [0, 1, 2])[0])
    d = (outs[0] - outs[1]).abs().flatten()
    assert float(d.median()) < 1e-5 and float((d < 1e-3).float().mean()) > 0.85, d
    # the batched gradient itself is exact: row r of the result is the gradient of the samples that used row r
    vq = PassBackend(n, L).vqc
    th = 0.3 * torch.randn(2, vq.n_params, generator=g)
    ang = torch.rand(6, n, generator=g) * math.pi
    cot = torch.randn(6, 1, generator=g)
    gb = vq.adjoint_grad(ang, th, cot)
    assert torch.allclose(gb[0], vq.adjoint_grad(ang[0::2], th[0], cot[0::2]), atol=1e-6)
    assert torch.allclose(gb[1], vq.adjoint_grad(ang[1::2], th[1], cot[1::2]), atol=1e-6)


def test_deep_circuits_on_few_qubits_are_cut_into_launchable_passes():
    """12 qubits x 16 layers fits ONE tile, so the scheduler returns a single pass of 396 gates - more than a launch of
    the pass kernel takes (160).  compile_passes cuts it into consecutive passes over the same tile; order and content of
    the gates are unchanged and every piece gets its own segment plan."""
    from qfedx_b200.ops.large import MAX_OPS_PER_PASS, ansatz_gates, compile_passes
    gates = ansatz_gates(12, 16)
    passes = compile_passes(12, gates, 12, 3)
    _check_schedule(12, gates, passes, 12, 3)
    assert len(passes) >= 3 and all(1 <= len(p.ops) <= MAX_OPS_PER_PASS for p in passes)
    assert all((p.tile_qubits == passes[0].tile_qubits).all() for p in passes)
    assert all(p.plan is not None for p in passes)
    assert sum(len(p.ops) for p in passes) == 12 + 16 * (12 + 12)
