"""Fused local-training kernel, fused evaluation and K7 aggregation vs torch fp32 references."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.fl import FederatedDataset, TorchLocalTrainer, VQCFusedTrainer
from qfedx_b200.fl import secagg
from qfedx_b200.models import VQC
from qfedx_b200.ops.aggregate import _reference, fedavg_aggregate

pytestmark = pytest.mark.gpu


def _clients(k, n_per, nq, classes, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.rand(n_per + 3 * i, nq, generator=g) * math.pi, torch.randint(0, classes, (n_per + 3 * i,), generator=g))
            for i in range(k)]


ADAM_EPS = 1e-4  # parameters with analytically zero gradient see only rounding noise; a visible eps keeps
                 # Adam from turning that noise (1e-9 in fp32, 1e-17 in fp64) into +-lr steps


def _torch_sequential_reference(model64, g0, clients, epochs, bs, lr, opt):
    """float64 eager loop over the same (unshuffled) minibatches."""
    from qfedx_b200.fl.trainers import make_optimizer
    rows = []
    for X, y in clients:
        model64.load_flat_params(g0.double())
        o = make_optimizer(model64.parameters(), opt, lr, 0.9, eps=ADAM_EPS)
        X = X.double()
        for _ in range(epochs):
            for lo in range(0, len(X), bs):
                o.zero_grad()
                model64.loss(model64(X[lo:lo + bs]), y[lo:lo + bs]).backward()
                o.step()
        rows.append(model64.flat_params().clone())
    return torch.stack(rows)


@pytest.mark.parametrize("lane_mode", ["latency", "throughput"])
@pytest.mark.parametrize("nq,classes,opt,cluster", [(4, 2, "adam", 1), (8, 2, "adam", 2), (6, 3, "sgd", 1),
                                                    (8, 3, "sgd", 4), (10, 2, "adam", 1), (1, 2, "sgd", 1),
                                                    (2, 2, "adam", 1)])
def test_fused_local_training_matches_torch_loop(nq, classes, opt, cluster, lane_mode):
    torch.manual_seed(0)
    clients = _clients(3, 40, nq, classes)
    model = VQC(nq, 2, classes)
    g0 = model.flat_params()
    gpu_data = FederatedDataset.from_clients(clients, "cuda")
    fused = VQCFusedTrainer(VQC(nq, 2, classes).cuda(), gpu_data, epochs=2, batch_size=16, lr=0.05, optimizer=opt,
                            cluster_size=cluster, shuffle=False, eps=ADAM_EPS,
                            lane_mode=lane_mode)                   # no shuffling: both walk the same minibatches
    out, _ = fused.train(g0.cuda(), [0, 1, 2], 0)
    exp = _torch_sequential_reference(model.double(), g0, clients, 2, 16, 0.05, opt)
    np.testing.assert_allclose(out.cpu().numpy(), exp.numpy(), atol=3e-4)


def test_tiered_cluster_sizes_match_the_sequential_reference():
    """Clients of very different sizes: the auto mode gives the longest client the widest cluster and the small ones
    narrower clusters (one launch per tier, side by side); every client's result still equals the float64 loop, in
    the caller's client order, and a wide cluster on a minibatch smaller than its slot count (idle warps) is exact."""
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(5)
    nq, classes = 8, 3
    clients = [(torch.rand(n, nq, generator=g) * math.pi, torch.randint(0, classes, (n,), generator=g))
               for n in (7, 150, 33, 64, 12)]
    model = VQC(nq, 2, classes)
    g0 = model.flat_params()
    data = FederatedDataset.from_clients(clients, "cuda")
    fused = VQCFusedTrainer(VQC(nq, 2, classes).cuda(), data, epochs=2, batch_size=16, lr=0.05, optimizer="adam",
                            shuffle=False, eps=ADAM_EPS)                # cluster_size=0: tiers
    order = [3, 0, 1, 4, 2]
    out, info = fused.train(g0.cuda(), order, 0)
    sizes = fused.last_cluster_sizes
    assert sizes[order.index(1)] == max(sizes) and min(sizes) < max(sizes), sizes
    exp = _torch_sequential_reference(model.double(), g0, clients, 2, 16, 0.05, "adam")
    np.testing.assert_allclose(out.cpu().numpy(), exp[order].numpy(), atol=3e-4)
    assert info["train_stats"].shape == (5, 2)
    wide = VQCFusedTrainer(VQC(nq, 2, classes).cuda(), data, epochs=2, batch_size=16, lr=0.05, optimizer="adam",
                           shuffle=False, eps=ADAM_EPS, cluster_size=8)   # 64 slots for 16 samples
    out8, _ = wide.train(g0.cuda(), order, 0)
    np.testing.assert_allclose(out8.cpu().numpy(), exp[order].numpy(), atol=3e-4)


def test_param_shift_training_equals_adjoint_training():
    clients = _clients(2, 24, 4, 2, seed=3)
    data = FederatedDataset.from_clients(clients, "cuda")
    m = VQC(4, 2, 2).cuda()
    a = VQCFusedTrainer(m, data, 1, 8, 0.05, "adam", grad_method="adjoint", shuffle=False, eps=ADAM_EPS)
    b = VQCFusedTrainer(m, data, 1, 8, 0.05, "adam", grad_method="param_shift", shuffle=False, cluster_size=2,
                        eps=ADAM_EPS)
    g0 = m.flat_params().cuda()
    oa, _ = a.train(g0, [0, 1], 0)
    ob, _ = b.train(g0, [0, 1], 0)
    np.testing.assert_allclose(oa.cpu().numpy(), ob.cpu().numpy(), atol=2e-4)


def test_fused_eval_matches_model():
    g = torch.Generator().manual_seed(1)
    X = (torch.rand(333, 8, generator=g) * math.pi).cuda()
    y = torch.randint(0, 3, (333,), generator=g).cuda()
    m = VQC(8, 3, 3).cuda()
    tr = VQCFusedTrainer(m, FederatedDataset.from_clients([(X, y)], "cuda"))
    loss, acc = tr.evaluate(m.flat_params().cuda(), X, y)
    z = m(X)
    assert loss == pytest.approx(float(torch.nn.functional.cross_entropy(z, y)), abs=1e-4)
    assert acc == pytest.approx(float((z.argmax(1) == y).float().mean()), abs=1e-6)


def test_shuffled_permutations_are_per_client():
    data = FederatedDataset.from_clients(_clients(4, 10, 4, 2), "cuda")
    perm = data.epoch_permutations(3).cpu().numpy()
    off = data.offsets.cpu().numpy()
    for e in range(3):
        for k in range(4):
            seg = perm[e, off[k]:off[k + 1]]
            assert sorted(seg.tolist()) == list(range(off[k], off[k + 1]))


@pytest.mark.parametrize("K,D", [(4, 96), (8, 113859), (64, 102), (5, 70000), (3, 4099)])
@pytest.mark.parametrize("mode,clip,secagg_on", [("weights", 0.0, False), ("delta", 0.0, False), ("delta", 0.7, False),
                                                 ("delta", 0.7, True)])
def test_fused_aggregation_matches_reference(K, D, mode, clip, secagg_on):
    g = torch.Generator().manual_seed(K + D)
    glob = torch.randn(D, generator=g)
    cp = glob + 0.02 * torch.randn(K, D, generator=g) * (1 + torch.arange(K)[:, None])
    w = torch.randint(1, 500, (K,), generator=g).float()
    part = (torch.rand(K, generator=g) > 0.3).int()
    part[0] = 1
    n_wrap = D // 2 if mode == "delta" else 0
    res = fedavg_aggregate(cp.cuda(), glob.cuda(), w.cuda(), part.cuda(), mode=mode, n_wrap=n_wrap, clip=clip,
                           server_lr=0.9, seed=11, rnd=4, secagg=secagg_on, want_norms=True, want_masked=secagg_on)
    exp = _reference(cp.double(), glob.double(), w, part, mode, n_wrap, clip, 0.0, 0.9, 11, 4, False, False)
    np.testing.assert_allclose(res.new_global.cpu().numpy(), exp.new_global.numpy(), atol=2e-5)
    if mode == "delta":
        np.testing.assert_allclose(res.norms.cpu().numpy(), exp.norms.numpy(), rtol=1e-4)
    if secagg_on and K * D < 10 ** 5:
        # masked messages carry the masks the host PRNG predicts and cancel in the sum
        idx = torch.nonzero(part).flatten().tolist()
        msgs = res.masked.cpu().numpy().view(np.uint64)[idx]
        host = secagg.mask_messages(np.zeros((K, D)), 11, 4, idx)[idx]
        with np.errstate(over="ignore"):
            payload = (msgs - host).view(np.int64).astype(np.float64) / secagg.FIXED_SCALE
        assert np.abs(payload).max() < 10.0                      # masks removed -> small fixed-point payloads
        assert np.abs(secagg.decode_fixed(msgs[0])).max() > 1e3  # a single message is hidden


def test_dp_noise_statistics():
    K, D, C, sigma = 8, 200000, 0.5, 1.3
    glob = torch.zeros(D).cuda()
    cp = torch.zeros(K, D).cuda()
    w = torch.ones(K).cuda()
    res = fedavg_aggregate(cp, glob, w, mode="delta", clip=C, sigma=sigma, seed=3, rnd=1)
    upd = res.new_global
    # mean of K independent N(0, (sigma C)^2) draws, each weighted 1/K
    assert float(upd.std()) == pytest.approx(sigma * C / math.sqrt(K), rel=0.02)
    assert abs(float(upd.mean())) < 5e-3
    res2 = fedavg_aggregate(cp, glob, w, mode="delta", clip=C, sigma=sigma, seed=3, rnd=1)
    assert torch.equal(res2.new_global, upd)                                 # keyed, reproducible
    res3 = fedavg_aggregate(cp, glob, w, mode="delta", clip=C, sigma=sigma, seed=3, rnd=2)
    assert not torch.equal(res3.new_global, upd)


def test_device_feistel_matches_host_and_fused_shuffle_uses_every_sample():
    import ctypes
    from qfedx_b200.fl import prng
    from qfedx_b200.ops import _lib
    for n, key in ((1106, 99), (33, 7), (4190, 2 ** 63 + 5)):
        out = torch.empty(n, dtype=torch.int32, device="cuda")
        _lib.call("qfx_debug_feistel", _lib.c_int(n), _lib.c_u64(key), _lib.ptr(out), _lib.stream_ptr())
        np.testing.assert_array_equal(out.cpu().numpy(), prng.feistel_permutation(n, key))
    # one epoch with batch = whole client: shuffled and unshuffled training must agree (order-independent mean)
    clients = _clients(2, 30, 4, 2, seed=5)
    data = FederatedDataset.from_clients(clients, "cuda")
    m = VQC(4, 1, 2).cuda()
    g0 = m.flat_params().cuda()
    a = VQCFusedTrainer(m, data, 1, 64, 0.05, "sgd", shuffle=False)
    b = VQCFusedTrainer(m, data, 1, 64, 0.05, "sgd", shuffle=True, shuffle_mode="feistel", seed=3)
    oa, _ = a.train(g0, [0, 1], 0)
    ob, _ = b.train(g0, [0, 1], 0)
    np.testing.assert_allclose(oa.cpu().numpy(), ob.cpu().numpy(), atol=1e-5)
