"""K5 pass engine and K6 fused swap (loopback on one GPU) vs the torch reference."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.ops import ref
from qfedx_b200.ops import statevector as svop

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,L,ent,tile", [(13, 2, "cnot", 12), (11, 3, "cz", 8), (16, 1, "cnot", 12), (5, 2, "cnot", 4)])
def test_pass_engine_matches_reference(n, L, ent, tile):
    from qfedx_b200.ops.large import LargeVQC
    g = torch.Generator().manual_seed(n)
    B = 3
    ang = (torch.rand(B, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    vqc = LargeVQC(n, L, ent, tile_bits=tile)
    st = vqc.states(ang, theta.reshape(1, -1))
    st_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double().reshape(1, L, n, 2), None, "ry", ent)
    np.testing.assert_allclose(st.cpu().numpy(), st_ref.numpy(), atol=3e-5)
    np.testing.assert_allclose(vqc.expvals(ang, theta.reshape(1, -1)).cpu().numpy(),
                               ref.expval_z_all(st_ref).numpy(), atol=3e-5)


def test_param_shift_gradient_matches_adjoint_kernel():
    from qfedx_b200.ops.large import LargeVQC
    n, L, S = 10, 2, 4
    g = torch.Generator().manual_seed(1)
    ang = (torch.rand(S, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    cot = torch.randn(S, n, generator=g).cuda()
    grad = LargeVQC(n, L).param_shift_grad(ang, theta, cot, max_states=64)
    gadj, _ = svop.sv_backward(ang, theta.reshape(1, L, n, 2), cot)            # register-tiled adjoint (K4a)
    np.testing.assert_allclose(grad.cpu().numpy(), gadj.reshape(-1).cpu().numpy(), atol=3e-4)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_swap_fused_loopback(world):
    from qfedx_b200.ops.large import ansatz_gates
    from qfedx_b200.parallel.sharded import ShardedState, plan_sharded
    n, L = 15, 2
    g = torch.Generator().manual_seed(world)
    ang = (torch.rand(1, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    gates = ansatz_gates(n, L)
    gbits = int(math.log2(world))
    plan = plan_sharded(n, gbits, gates, tile_bits=10)
    assert plan.n_swaps >= 1
    st = ShardedState(n, world, loopback=True)
    st.reset_zero_state()
    st.run(plan, theta, ang)
    full = st.gather_logical(plan)
    st_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double().reshape(1, L, n, 2))[0]
    np.testing.assert_allclose(full.cpu().numpy(), st_ref.numpy(), atol=3e-5)
    np.testing.assert_allclose(st.expvals(plan).cpu().numpy(), ref.expval_z_all(st_ref[None])[0].numpy(), atol=3e-5)


def test_large_vqc_trainer_param_shift_and_spsa_learn():
    """Config-4 style pipeline on the pass engine: parameter-shift gradients, secure aggregation."""
    from qfedx_b200.fl import FederatedDataset, FederatedRunner, FLConfig
    from qfedx_b200.fl.trainers_large import LargeVQCTrainer, PassBackend
    from qfedx_b200.models import VQC
    torch.manual_seed(0)
    n = 12
    g = torch.Generator().manual_seed(0)
    def client(seed):
        gg = torch.Generator().manual_seed(seed)
        y = torch.randint(0, 2, (24,), generator=gg)
        x = torch.rand(24, n, generator=gg) * 0.3 + y[:, None] * 2.0     # separable in every angle
        return x, y
    clients = [client(s) for s in range(3)]
    data = FederatedDataset.from_clients(clients, "cuda")
    model = VQC(n, 1, 2).cuda()
    tr = LargeVQCTrainer(model, data, PassBackend(n, 1), epochs=1, batch_size=8, lr=0.2, grad_method="param_shift")
    Xt = torch.cat([c[0] for c in clients]).cuda()
    yt = torch.cat([c[1] for c in clients]).cuda()
    run = FederatedRunner(tr, FLConfig(num_rounds=3, mode="delta", secagg=True, clip=5.0, quiet=True), (Xt, yt))
    out = run.run()
    assert out["history"]["loss"][-1] < out["history"]["loss"][0]
    # parameter-shift gradient of the trainer's backend equals the adjoint kernel's on a 10-qubit check
    tr2 = LargeVQCTrainer(model, data, PassBackend(n, 1), epochs=1, batch_size=8, lr=0.2, grad_method="spsa")
    w, _ = tr2._train_one(tr2.global_flat(), 0)
    assert torch.isfinite(w).all()


def test_sharded_backend_loopback_trainer_runs():
    from qfedx_b200.fl import FederatedDataset
    from qfedx_b200.fl.trainers_large import LargeVQCTrainer, PassBackend, ShardedBackend
    from qfedx_b200.models import VQC
    n = 12
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(4, n, generator=g) * math.pi).cuda()
    theta = torch.randn(2 * n, generator=g).cuda()
    a = ShardedBackend(n, 1, 4, 0, torch.device("cuda"), loopback=True).expvals(x, theta)
    b = PassBackend(n, 1).expvals(x, theta)
    np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), atol=3e-5)


def _parity_reference(st, n_local, masks, high):
    idx = (high << n_local) | np.arange(1 << n_local, dtype=np.int64)
    p = (st.abs() ** 2).cpu().double().numpy()
    out = np.zeros((st.shape[0], len(masks)))
    for j, m in enumerate(masks):
        par = np.zeros_like(idx)
        v = idx & m
        while v.any():
            par ^= v & 1
            v >>= 1
        out[:, j] = (p * (1 - 2 * par)).sum(1)
    return out


@pytest.mark.parametrize("n_local,high", [(1, 3), (2, 0), (5, 6), (8, 1), (9, 2), (10, 7), (11, 5), (12, 0), (14, 5), (17, 3)])
def test_parity_readout_kernel_matches_torch(n_local, high):
    """qfx_sv_large_expz_masks on random states, arbitrary masks, with rank bits above the shard.  The shard sizes walk
    through every index-bit class of the kernel: the pair bit, the thread's bits, the folded bits 9 and 10, the table
    bytes, and rank bits landing in each of them."""
    from qfedx_b200.ops.large import expvals_parity, expvals_z
    B = 3
    g = torch.Generator().manual_seed(7 + n_local)
    st = torch.complex(torch.randn(B, 1 << n_local, generator=g), torch.randn(B, 1 << n_local, generator=g)).cuda()
    masks = [int(m) for m in torch.randint(1, 1 << (n_local + 3), (9,), generator=g)] + [1, 1 << (n_local - 1), (1 << (n_local + 3)) - 1]
    got = expvals_parity(st.to(torch.complex64), n_local, masks, high_bits=high).cpu().double().numpy()
    want = _parity_reference(st, n_local, masks, high)
    scale = float((st.abs() ** 2).sum(1).max())
    np.testing.assert_allclose(got, want, atol=2e-5 * scale)
    # <Z_q> of every qubit, rank bits included, is the same kernel with one-bit masks
    n_total = n_local + 3
    gz = expvals_z(st.to(torch.complex64), n_local, n_total, high_bits=high).cpu().double().numpy()
    np.testing.assert_allclose(gz, _parity_reference(st, n_local, [1 << q for q in range(n_total)], high), atol=2e-5 * scale)


@pytest.mark.parametrize("world", [2, 8])
def test_sharded_parity_readout_replaces_final_ring(world):
    """Two layers: the first ring is simulated (swaps), the second only changes the readout masks."""
    from qfedx_b200.fl.trainers_large import ShardedBackend
    n, L = 14, 2
    g = torch.Generator().manual_seed(world)
    ang = (torch.rand(2, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    be = ShardedBackend(n, L, world, 0, torch.device("cuda"), loopback=True)
    assert be.plan.n_swaps >= 1
    st_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double().reshape(1, L, n, 2))
    np.testing.assert_allclose(be.expvals(ang, theta).cpu().numpy(), ref.expval_z_all(st_ref).numpy(), atol=3e-5)
    be.n_out = 3
    np.testing.assert_allclose(be.expvals(ang, theta).cpu().numpy(), ref.expval_z_all(st_ref).numpy()[:, :3], atol=3e-5)


def test_param_shift_prefix_sharing_multi_pass():
    """Shifted angles that first appear in a later pass start from the shared unshifted prefix; the last layer's RZ
    angles are skipped (zero gradient).  Checked against the adjoint kernel (small tiles force several passes) and
    against explicit shifts of the full circuit on the 12-qubit-tile engine."""
    from qfedx_b200.ops.large import LargeVQC
    n, L, S = 10, 2, 3
    g = torch.Generator().manual_seed(5)
    ang = (torch.rand(S, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    cot = torch.randn(S, n, generator=g).cuda()
    vqc = LargeVQC(n, L, tile_bits=6)
    assert len(vqc.readout_passes) >= 3 and len(set(vqc._first_pass_of_param().values())) >= 3
    grad = vqc.param_shift_grad(ang, theta, cot, max_states=32)
    gadj, _ = svop.sv_backward(ang, theta.reshape(1, L, n, 2), cot)
    np.testing.assert_allclose(grad.cpu().numpy(), gadj.reshape(-1).cpu().numpy(), atol=3e-4)

    n, L, S = 14, 2, 2
    ang = (torch.rand(S, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    cot = torch.randn(S, 3, generator=g).cuda()                      # three logit qubits read out
    vqc = LargeVQC(n, L)
    fp = vqc._first_pass_of_param()
    assert len(vqc.readout_passes) >= 2 and max(fp.values()) >= 1
    grad = vqc.param_shift_grad(ang, theta, cot, max_states=64)
    later = [p for p, k in fp.items() if k >= 1][:3]
    for p in [0, 5, 2 * n + 4, 2 * n + 1] + later:                   # incl. a last-layer RZ angle (index 2n+1)
        tp, tm = theta.clone(), theta.clone()
        tp[p] += math.pi / 2
        tm[p] -= math.pi / 2
        d = vqc.expvals(ang, tp.reshape(1, -1), 3) - vqc.expvals(ang, tm.reshape(1, -1), 3)
        assert float(grad[p]) == pytest.approx(0.5 * float((d * cot).sum()), abs=3e-4)
    assert float(grad[2 * n + 1]) == 0.0


@pytest.mark.parametrize("n,L,S,n_out,tile,ent", [(10, 2, 4, 10, 12, "cnot"), (10, 3, 3, 2, 6, "cnot"), (14, 2, 2, 3, 12, "cnot"),
                                                (13, 2, 2, 1, 12, "cz"), (20, 2, 2, 1, 12, "cnot")])
def test_adjoint_on_the_pass_engine_matches_the_other_gradients(n, L, S, n_out, tile, ent):
    """One forward + one backward sweep over the compiled passes gives the gradient of the shift rule (and, where a
    register kernel exists, of K4a)."""
    from qfedx_b200.ops.large import LargeVQC
    g = torch.Generator().manual_seed(100 + n)
    ang = (torch.rand(S, n, generator=g) * 0.8).cuda()              # states near |0..0>: readouts far from zero, so the
    theta = (0.3 * torch.randn(L * n * 2, generator=g)).cuda()      # gradients sit well above fp32 rounding noise
    cot = torch.randn(S, n_out, generator=g).cuda()
    vqc = LargeVQC(n, L, entangler=ent, tile_bits=tile)
    gadj = vqc.adjoint_grad(ang, theta, cot)
    gshift = vqc.param_shift_grad(ang, theta, cot, max_states=64)
    scale = float(gshift.abs().max())
    assert scale > 1e-2
    # relative to the largest component: both are fp32 sums over up to 2^20 amplitudes
    assert float((gadj - gshift).abs().max()) < 2e-3 * scale + 1e-5, (gadj, gshift)
    if n <= 10:
        full = torch.zeros(S, n, device="cuda")
        full[:, :n_out] = cot
        greg, _ = svop.sv_backward(ang, theta.reshape(1, L, n, 2), full, entangler=ent)
        assert float((gadj - greg.reshape(-1)).abs().max()) < 2e-3 * scale + 1e-5


def test_large_vqc_trainer_with_adjoint_gradients_matches_param_shift_training():
    from qfedx_b200.fl import FederatedDataset
    from qfedx_b200.fl.trainers_large import LargeVQCTrainer, PassBackend
    from qfedx_b200.models import VQC
    n, L = 12, 2
    g = torch.Generator().manual_seed(3)
    clients = [((torch.rand(6, n, generator=g) * math.pi), torch.randint(0, 2, (6,), generator=g)) for _ in range(2)]
    data = FederatedDataset.from_clients(clients, "cuda")
    torch.manual_seed(0)
    model = VQC(n, L, 2).cuda()
    outs = {}
    for method in ("param_shift", "adjoint"):
        gen = torch.Generator(device="cuda").manual_seed(7)
        tr = LargeVQCTrainer(model, data, PassBackend(n, L), epochs=1, batch_size=3, lr=0.05, grad_method=method, generator=gen)
        outs[method], _ = tr.train(tr.global_flat(), [0, 1], 0)
    # Adam turns gradients that are zero up to rounding into lr-sized steps of either sign, so a few coordinates may
    # differ by ~lr between two correct gradient methods; everything else must agree closely
    diff = (outs["adjoint"] - outs["param_shift"]).abs().flatten()
    assert float((diff < 2e-3).float().mean()) >= 0.95 and float(diff.median()) < 1e-4, diff


def test_lockstep_adjoint_trainer_on_the_gpu_matches_per_client_training():
    from qfedx_b200.fl import FederatedDataset
    from qfedx_b200.fl.trainers_large import LargeVQCTrainer, PassBackend
    from qfedx_b200.models import VQC
    n, L = 13, 2
    g = torch.Generator().manual_seed(4)
    clients = [((torch.rand(5 + k, n, generator=g) * 0.8), torch.randint(0, 2, (5 + k,), generator=g)) for k in range(3)]
    data = FederatedDataset.from_clients(clients, "cuda")
    torch.manual_seed(0)
    model = VQC(n, L, 2).cuda()
    outs = []
    for mode in ("seq", "lock"):
        gen = torch.Generator(device="cuda").manual_seed(7)
        tr = LargeVQCTrainer(model, data, PassBackend(n, L), epochs=2, batch_size=2, lr=0.05, grad_method="adjoint", generator=gen)
        if mode == "seq":
            outs.append(torch.stack([tr._train_one(tr.global_flat(), k)[0] for k in range(3)]))
        else:
            outs.append(tr.train(tr.global_flat(), [0, 1, 2], 0)[0])
    d = (outs[0] - outs[1]).abs().flatten()
    assert float(d.median()) < 1e-4 and float((d < 2e-3).float().mean()) > 0.85, d
