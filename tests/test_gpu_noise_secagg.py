"""Round-2 GPU features: readout noise in the K3 epilogue of the training / evaluation kernels, and the aggregation
kernel keyed by per-pair secrets from the key agreement."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.fl import FederatedDataset, VQCFusedTrainer
from qfedx_b200.fl.secagg import KeyAgreement
from qfedx_b200.models import VQC
from qfedx_b200.noise import NoiseConfig, apply_to_expvals, noisy_expvals_cuda, readout_affine
from qfedx_b200.ops.aggregate import _reference, fedavg_aggregate

pytestmark = pytest.mark.gpu
ADAM_EPS = 1e-4


def test_readout_noise_pass_matches_apply_to_expvals():
    g = torch.Generator().manual_seed(0)
    E = (torch.rand(4096, 3, generator=g) * 2 - 1).cuda()
    cfg = NoiseConfig("depolarizing", p=0.1, readout_p01=0.03, readout_p10=0.06)
    assert torch.allclose(noisy_expvals_cuda(E, cfg), apply_to_expvals(E, cfg), atol=1e-6)     # affine part: exact
    amp = NoiseConfig("amplitude", gamma=0.2)
    assert torch.allclose(noisy_expvals_cuda(E, amp), apply_to_expvals(E, amp), atol=1e-6)
    # shots: same distribution as the torch restatement - mean of the affine value, variance (1 - E'^2) / shots;
    # 64 shots take the exact Bernoulli path of the kernel, 2000 the normal approximation
    for shots in (64, 2000):
        c = NoiseConfig("depolarizing", p=0.1, shots=shots)
        a, b = readout_affine(c)
        const = torch.full((20000, 1), 0.4, device="cuda")
        s1 = noisy_expvals_cuda(const, c, key=1)
        s2 = noisy_expvals_cuda(const, c, key=2)
        target = a * 0.4 + b
        assert float(s1.mean()) == pytest.approx(target, abs=4 * math.sqrt((1 - target ** 2) / shots / 20000))
        assert float(s1.std()) == pytest.approx(math.sqrt((1 - target ** 2) / shots), rel=0.05)
        assert not torch.equal(s1, s2)                                   # keyed: another key, another sample
        assert torch.equal(s1, noisy_expvals_cuda(const, c, key=1))      # ... and reproducible
        k = (s1 + 1) * shots / 2                                         # outcomes are whole shot counts
        assert torch.allclose(k, k.round(), atol=1e-3)


def _clients(k, n_per, nq, classes, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.rand(n_per + 3 * i, nq, generator=g) * math.pi, torch.randint(0, classes, (n_per + 3 * i,), generator=g))
            for i in range(k)]


@pytest.mark.parametrize("nq,classes,grad", [(4, 2, "adjoint"), (6, 3, "adjoint"), (4, 2, "param_shift")])
def test_fused_training_and_eval_with_affine_readout_noise_match_the_torch_model(nq, classes, grad):
    """Depolarizing + confusion-matrix noise is deterministic (affine on <Z>): the fused kernels must reproduce the
    float64 eager loop through VQC.forward, which applies noise.apply_to_expvals."""
    from qfedx_b200.fl.trainers import make_optimizer
    noise = NoiseConfig("depolarizing", p=0.15, readout_p01=0.02, readout_p10=0.05)
    torch.manual_seed(0)
    clients = _clients(2, 32, nq, classes)
    ref = VQC(nq, 2, classes, noise=noise).double()
    g0 = ref.flat_params().float()
    data = FederatedDataset.from_clients(clients, "cuda")
    fused = VQCFusedTrainer(VQC(nq, 2, classes).cuda(), data, epochs=2, batch_size=16, lr=0.05, optimizer="adam",
                            shuffle=False, eps=ADAM_EPS, grad_method=grad, noise=noise)
    out, _ = fused.train(g0.cuda(), [0, 1], 0)
    rows = []
    for X, y in clients:
        ref.load_flat_params(g0.double())
        opt = make_optimizer(ref.parameters(), "adam", 0.05, 0.9, eps=ADAM_EPS)
        for _ in range(2):
            for lo in range(0, len(X), 16):
                opt.zero_grad()
                ref.loss(ref(X[lo:lo + 16].double()), y[lo:lo + 16]).backward()
                opt.step()
        rows.append(ref.flat_params().clone())
    np.testing.assert_allclose(out.cpu().numpy(), torch.stack(rows).numpy(), atol=3e-4)
    # and differs from the noise-free training
    clean = VQCFusedTrainer(VQC(nq, 2, classes).cuda(), data, epochs=2, batch_size=16, lr=0.05, optimizer="adam",
                            shuffle=False, eps=ADAM_EPS, grad_method=grad)
    out0, _ = clean.train(g0.cuda(), [0, 1], 0)
    assert float((out - out0).abs().max()) > 1e-3
    # evaluation through the same epilogue
    X, y = clients[0]
    loss, acc = fused.evaluate(g0.cuda(), X.cuda(), y.cuda())
    ref.load_flat_params(g0.double())
    with torch.no_grad():
        z = ref(X.double())
    assert loss == pytest.approx(float(ref.loss(z, y)), abs=2e-5)


def test_shot_noise_reaches_the_fused_kernels():
    noise = NoiseConfig(shots=16)
    clients = _clients(2, 64, 4, 2, seed=2)
    data = FederatedDataset.from_clients(clients, "cuda")
    m = VQC(4, 2, 2).cuda()
    g0 = m.flat_params().cuda()
    exact = VQCFusedTrainer(m, data, 1, 16, 0.05, shuffle=False)
    shot = VQCFusedTrainer(VQC(4, 2, 2).cuda(), data, 1, 16, 0.05, shuffle=False, noise=noise)
    a, _ = exact.train(g0, [0, 1], 0)
    b, _ = shot.train(g0, [0, 1], 0)
    c, _ = shot.train(g0, [0, 1], 0)
    d, _ = shot.train(g0, [0, 1], 1)
    assert torch.isfinite(b).all() and not torch.allclose(a, b, atol=1e-5)
    assert torch.equal(b, c)                        # keyed by (seed, round): the same round reproduces
    assert not torch.equal(b, d)                    # another round draws other shots
    X, y = clients[0]
    l0, _ = exact.evaluate(g0, X.cuda(), y.cuda())
    l1, _ = shot.evaluate(g0, X.cuda(), y.cuda())
    assert math.isfinite(l1) and l1 != l0


@pytest.mark.parametrize("K,D", [(6, 102), (5, 4096)])
def test_aggregation_with_pair_secrets_matches_the_host_protocol(K, D):
    g = torch.Generator().manual_seed(K + D)
    glob = torch.randn(D, generator=g)
    cp = glob + 0.05 * torch.randn(K, D, generator=g)
    w = torch.randint(1, 50, (K,), generator=g).float()
    part = torch.ones(K, dtype=torch.int32)
    part[1] = 0
    keys = KeyAgreement(K, seed=3).table()
    kw = dict(mode="delta", n_wrap=D // 2, clip=0.5, server_lr=0.9, seed=7, rnd=2, secagg=True)
    dev = fedavg_aggregate(cp.cuda(), glob.cuda(), w, part, want_masked=True, pair_keys=keys, **kw)
    host = _reference(cp, glob, w, part, kw["mode"], kw["n_wrap"], kw["clip"], 0.0, kw["server_lr"], kw["seed"], kw["rnd"],
                      True, True, pair_keys=keys)
    assert torch.allclose(dev.new_global.cpu(), host.new_global, atol=2e-6)
    # the masked messages themselves: bit-identical fixed-point words for the participating rows wherever the float
    # -> fixed conversion agrees (the kernel multiplies in fp32, the host in fp64: compare the unmasked sum instead)
    seedmask = fedavg_aggregate(cp.cuda(), glob.cuda(), w, part, want_masked=True, **kw)
    assert not torch.equal(dev.masked, seedmask.masked)              # other secrets, other masks ...
    assert torch.allclose(dev.new_global, seedmask.new_global, atol=1e-6)   # ... same aggregate
    # mask words of one pair follow the shared stream: message difference between two key tables is the mask difference
    assert int((dev.masked[0] != seedmask.masked[0]).sum()) > D // 2
