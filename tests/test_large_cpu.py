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
