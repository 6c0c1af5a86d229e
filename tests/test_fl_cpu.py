"""FL semantics on CPU: the ROADMAP's verification checks as tests."""
import math
import random

import numpy as np
import pytest
import torch

from qfedx_b200 import set_seeds
from qfedx_b200.data import partition, synthetic_mnist
from qfedx_b200.fl import (FederatedDataset, FederatedRunner, FLConfig, RDPAccountant, TorchLocalTrainer,
                           VQCFusedTrainer, federated_learning, gaussian_epsilon_closed_form, inject_dropouts,
                           quantum_federated_learning, rdp_sampled_gaussian, sample_clients)
from qfedx_b200.fl import prng, secagg
from qfedx_b200.fl.dp import clip_scale, wrap_angles
from qfedx_b200.models import VQC, TinyCNN, flatten_state
from qfedx_b200.ops.aggregate import fedavg_aggregate
from qfedx_b200.oracle import fedavg as ofa


def test_aggregator_reconstructs_weighted_mean():
    g = torch.Generator().manual_seed(0)
    cp = torch.randn(5, 37, generator=g)
    w = torch.tensor([10., 20., 5., 1., 64.])
    res = fedavg_aggregate(cp, torch.zeros(37), w, mode="weights")
    sd = [({"p": cp[k].numpy()}, float(w[k])) for k in range(5)]
    np.testing.assert_allclose(res.new_global.numpy(), ofa.fedavg_state_dicts(sd)["p"], atol=1e-6)


def test_delta_mode_equals_weights_mode_without_privacy():
    g = torch.Generator().manual_seed(1)
    glob = torch.randn(20, generator=g)
    cp = glob + 0.1 * torch.randn(4, 20, generator=g)
    w = torch.tensor([1., 2., 3., 4.])
    a = fedavg_aggregate(cp, glob, w, mode="weights").new_global
    b = fedavg_aggregate(cp, glob, w, mode="delta").new_global
    assert torch.allclose(a, b, atol=1e-6)


def test_clip_and_participation_match_oracle():
    g = torch.Generator().manual_seed(2)
    glob = torch.randn(16, generator=g)
    cp = glob + torch.randn(6, 16, generator=g)
    w = torch.arange(1., 7.)
    part = torch.tensor([1, 0, 1, 1, 0, 1])
    res = fedavg_aggregate(cp, glob, w, part, mode="delta", clip=0.5)
    exp = ofa.aggregate_flat((cp - glob).numpy(), w.numpy(), clip=0.5, participate=part.numpy())
    np.testing.assert_allclose((res.new_global - glob).numpy(), exp, atol=1e-6)
    assert torch.all(res.norms > 0)


def test_angle_wrapping():
    d = torch.tensor([[3.5, -3.5, 0.2, 7.0]])
    out = wrap_angles(d, 3)
    np.testing.assert_allclose(out[0, :3].numpy(), ofa.wrap_angle(d[0, :3].numpy()), atol=1e-6)
    assert out[0, 3] == 7.0


def test_masked_sum_equals_raw_sum_exactly():
    rng = np.random.default_rng(0)
    vals = rng.normal(size=(6, 50))
    msgs = secagg.mask_messages(vals, seed=7, rnd=3)
    # a single masked message looks nothing like the value it hides
    assert np.abs(secagg.decode_fixed(msgs[0]) - vals[0]).max() > 1e3
    with np.errstate(over="ignore"):
        raw = secagg.encode_fixed(vals).sum(axis=0, dtype=np.uint64)
    assert np.array_equal(msgs.sum(axis=0, dtype=np.uint64), raw)        # bit-exact cancellation
    np.testing.assert_allclose(secagg.unmask_sum(msgs), vals.sum(0), atol=1e-8)


def test_dropout_repair():
    rng = np.random.default_rng(1)
    vals = rng.normal(size=(5, 12))
    msgs = secagg.mask_messages(vals, seed=1, rnd=0)                      # all five masked against each other
    survivors, dropped = [0, 1, 3], [2, 4]
    with np.errstate(over="ignore"):
        tot = msgs[survivors].sum(axis=0, dtype=np.uint64)
    fixed = secagg.unmask_dropped(tot, 1, 0, survivors, dropped, 5)
    np.testing.assert_allclose(secagg.decode_fixed(fixed), vals[survivors].sum(0), atol=1e-8)


def test_secagg_aggregate_matches_plain():
    g = torch.Generator().manual_seed(3)
    glob = torch.randn(24, generator=g)
    cp = glob + 0.3 * torch.randn(5, 24, generator=g)
    w = torch.tensor([3., 1., 4., 1., 5.])
    a = fedavg_aggregate(cp, glob, w, mode="delta", clip=1.0).new_global
    b = fedavg_aggregate(cp, glob, w, mode="delta", clip=1.0, secagg=True, seed=9, rnd=2).new_global
    assert torch.allclose(a, b, atol=1e-6)


def test_prng_is_a_pure_function():
    a = prng.pair_mask(5, 1, 2, 4, 8, 16)
    b = prng.pair_mask(5, 1, 4, 2, 8, 16)
    assert np.array_equal(a, b) and a.dtype == np.uint64
    assert not np.array_equal(a, prng.pair_mask(5, 2, 2, 4, 8, 16))


def test_accountant_closed_form_and_monotonicity():
    acc = RDPAccountant(orders=tuple(range(2, 2000)))
    acc.step(1.0, 2.0, 10)
    eps = acc.epsilon(1e-5, improved=False)
    assert eps == pytest.approx(gaussian_epsilon_closed_form(2.0, 10, 1e-5), rel=2e-3)
    assert acc.epsilon(1e-5, improved=True) <= eps
    # subsampling amplifies privacy; more noise is more private; more rounds cost more
    assert rdp_sampled_gaussian(0.1, 1.0, 8) < rdp_sampled_gaussian(1.0, 1.0, 8)
    e = []
    for sigma in (0.8, 1.5, 3.0):
        a = RDPAccountant(); a.step(0.2, sigma, 100); e.append(a.epsilon(1e-5))
    assert e[0] > e[1] > e[2] > 0
    a = RDPAccountant(); a.step(0.2, 1.0, 10); e10 = a.epsilon(1e-5); a.step(0.2, 1.0, 90)
    assert a.epsilon(1e-5) > e10
    b = RDPAccountant(); b.load_state_dict(a.state_dict())
    assert b.epsilon(1e-5) == pytest.approx(a.epsilon(1e-5))


def test_rdp_small_q_limit():
    # for small q the sampled-Gaussian RDP behaves like 2 q^2 alpha / sigma^2 (up to constants): O(q^2)
    r1, r2 = rdp_sampled_gaussian(0.01, 2.0, 4), rdp_sampled_gaussian(0.02, 2.0, 4)
    assert 3.0 < r2 / r1 < 5.0


def test_sampler_deterministic_and_sized():
    s1, s2 = sample_clients(64, 0.25, 3, seed=1), sample_clients(64, 0.25, 3, seed=1)
    assert np.array_equal(s1, s2) and len(s1) == 16 and len(set(s1.tolist())) == 16
    assert not np.array_equal(s1, sample_clients(64, 0.25, 4, seed=1))
    assert len(sample_clients(10, 1.0, 0)) == 10
    surv = inject_dropouts(np.arange(64), 0.5, 0, seed=1)
    assert 10 < len(surv) < 54
    assert len(inject_dropouts(np.arange(3), 1.0, 0)) == 1              # never empty


def _toy_clients(n=256, k=4, classes=2, seed=0):
    X, y = synthetic_mnist(n, classes, seed=seed)
    random.seed(seed)
    return partition(X, y, k, "iid")


def test_vqc_federated_loss_decreases_and_beats_chance():
    set_seeds(0)
    clients = _toy_clients(320, 4)
    Xt, yt = synthetic_mnist(200, 2, seed=5)
    out = quantum_federated_learning(clients, (Xt, yt), n_qubits=4, n_layers=2, num_classes=2, num_rounds=6,
                                     local_epochs=2, learning_rate=0.1, batch_size=32, quiet=True)
    assert out["history"]["loss"][-1] < out["history"]["loss"][0]
    assert out["accuracies"][-1] > 0.6


def test_tinycnn_federated_improves_over_round0():
    set_seeds(0)
    clients = _toy_clients(240, 3, classes=3)
    Xt, yt = synthetic_mnist(120, 3, seed=9)
    out = federated_learning(clients, (Xt, yt), num_rounds=2, local_epochs=1, learning_rate=0.05, batch_size=32,
                             num_classes=3, device="cpu", quiet=True)
    assert out["accuracies"][-1] > out["accuracies"][0]
    assert set(out["model"].state_dict().keys()) == {"conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
                                                     "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"}
    assert flatten_state(out["model"].state_dict()).numel() == 113859


def test_dp_secagg_dropout_run_and_checkpoint_resume(tmp_path):
    set_seeds(1)
    clients = [(torch.rand(40, 4) * math.pi, torch.randint(0, 2, (40,))) for _ in range(5)]
    data = FederatedDataset.from_clients(clients)
    cfg = FLConfig(num_rounds=4, mode="delta", clip=0.5, sigma=0.8, secagg=True, drop_prob=0.3, client_fraction=0.8,
                   checkpoint_every=2, checkpoint_dir=str(tmp_path), quiet=True, log_path=str(tmp_path / "log.jsonl"))
    tr = VQCFusedTrainer(VQC(4, 1, 2), data, epochs=1, batch_size=16, lr=0.05)
    out = FederatedRunner(tr, cfg, (clients[0][0], clients[0][1])).run()
    assert len(out["history"]["epsilon"]) == 4 and out["history"]["epsilon"][-1] > out["history"]["epsilon"][0]
    assert (tmp_path / "round_000004.pt").exists() and (tmp_path / "log.jsonl").read_text().count("\n") == 5
    cfg2 = FLConfig(**{**cfg.__dict__, "num_rounds": 6, "resume": True})
    tr2 = VQCFusedTrainer(VQC(4, 1, 2), data, epochs=1, batch_size=16, lr=0.05)
    r2 = FederatedRunner(tr2, cfg2, (clients[0][0], clients[0][1]))
    assert r2.round == 4 and torch.allclose(r2.global_flat, out["global_flat"])
    r2.run()
    assert r2.round == 6


def test_clip_scale():
    d = torch.tensor([[3., 4.], [0.3, 0.4]])
    np.testing.assert_allclose(clip_scale(d, 1.0).numpy(), [0.2, 1.0], atol=1e-7)


def test_spsa_trainer_runs():
    set_seeds(0)
    clients = [(torch.rand(32, 3) * math.pi, torch.randint(0, 2, (32,))) for _ in range(2)]
    data = FederatedDataset.from_clients(clients)
    m = VQC(3, 1, 2)
    tr = TorchLocalTrainer(m, data, epochs=1, batch_size=16, lr=0.05, optimizer="adam", grad_estimator="spsa")
    out, stats = tr.train(tr.global_flat(), [0, 1], 0)
    assert out.shape == (2, tr.dim) and not torch.allclose(out[0], tr.global_flat())


def test_feistel_permutation_is_a_keyed_bijection():
    for n in (1, 2, 3, 7, 32, 33, 100, 1106, 4190):
        p = prng.feistel_permutation(n, 12345)
        assert sorted(p.tolist()) == list(range(n))
    a, b = prng.feistel_permutation(500, 1), prng.feistel_permutation(500, 2)
    assert not np.array_equal(a, b)
    assert np.abs(np.corrcoef(np.arange(500), a)[0, 1]) < 0.2          # looks shuffled
