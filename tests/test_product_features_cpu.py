"""Product features of round 2 on CPU: noise in the training path, Iris, the centralized baseline, the sweep runner,
AUC, communication totals, the key-agreement secure aggregation with late drop-outs, and the DP accounting status."""
import json
import math
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from qfedx_b200.cli import run_experiment
from qfedx_b200.config import load_config
from qfedx_b200.fl import FederatedDataset, FederatedRunner, FLConfig, TorchLocalTrainer
from qfedx_b200.fl.secagg import KeyAgreement, mask_messages, unmask_dropped, decode_fixed
from qfedx_b200.models import VQC
from qfedx_b200.noise import NoiseConfig, apply_to_expvals, readout_affine
from qfedx_b200.ops.aggregate import fedavg_aggregate
from qfedx_b200.sweep import format_table, parse_grid, run_sweep
from qfedx_b200.utils.metrics import auc_from_logits, roc_auc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["device=cpu", "fl.quiet=true", "fl.num_rounds=3", "train.local_epochs=1"]


def _tmp(tmp_path):
    return [f"data.raw_folder={tmp_path}", f"data.processed_folder={tmp_path / 'p'}", "data.synthetic_scale=0.02"]


def test_readout_affine_matches_apply_to_expvals():
    E = torch.linspace(-1, 1, 9)[None]
    for cfg in (NoiseConfig("depolarizing", p=0.2), NoiseConfig("amplitude", gamma=0.3),
                NoiseConfig("depolarizing", p=0.1, readout_p01=0.04, readout_p10=0.07),
                NoiseConfig(readout_p01=0.1)):
        a, b = readout_affine(cfg)
        assert torch.allclose(apply_to_expvals(E, cfg), a * E + b, atol=1e-6)


def test_noise_config_reaches_the_model_and_degrades_the_signal(tmp_path):
    base = load_config("vqc4", SMALL + _tmp(tmp_path) + ["data.reducer=block"])
    noisy = load_config("vqc4_noisy", SMALL + _tmp(tmp_path) + ["data.reducer=block", "noise.p=0.5", "noise.shots=8"])
    a, b = run_experiment(base), run_experiment(noisy)
    assert a["final_loss"] is not None and b["final_loss"] is not None
    # the noisy run really took another path (the config reached model and trainer) and stayed finite
    assert math.isfinite(b["final_loss"]) and b["final_loss"] != a["final_loss"]
    # gradient through the affine part: a depolarized model's theta gradient is (1 - p) times the exact one
    torch.manual_seed(0)
    x = torch.rand(5, 4) * math.pi
    y = torch.randint(0, 2, (5,))
    m0, m1 = VQC(4, 2, 2), VQC(4, 2, 2, noise=NoiseConfig("depolarizing", p=0.25))
    m1.load_state_dict(m0.state_dict())
    E0 = m0.expvals(x)[:, :1]
    g0 = torch.autograd.grad(E0.sum(), m0.theta)[0]
    E1 = apply_to_expvals(m1.expvals(x)[:, :1], m1.noise)
    g1 = torch.autograd.grad(E1.sum(), m1.theta)[0]
    assert torch.allclose(g1, 0.75 * g0, atol=1e-6)


def test_iris_preset_learns(tmp_path):
    out = run_experiment(load_config("iris4", ["device=cpu", "fl.quiet=true", "fl.num_rounds=8", "train.local_epochs=3"]))
    assert out["clients"] == 4 and out["accuracies"][-1] > max(out["accuracies"][0], 0.5)
    assert 0.5 < out["final_auc"] <= 1.0 and out["comm_mb_total"] > 0


def test_centralized_baseline_pools_the_clients(tmp_path):
    out = run_experiment(load_config("iris4", ["device=cpu", "fl.quiet=true", "fl.num_rounds=4", "train.local_epochs=2",
                                               "train.centralized=true"]))
    assert out["centralized"] and out["clients"] == 1 and out["accuracies"][-1] > 0.5


def test_sweep_aggregates_over_seeds_and_arms():
    assert parse_grid(["fl.sigma=0,0.5", "data.alpha=0.1"]) == {"fl.sigma": [0, 0.5], "data.alpha": [0.1]}
    rows = run_sweep("iris4", ["device=cpu", "fl.num_rounds=2", "train.local_epochs=1", "fl.clip=1.0"],
                     ["fl.sigma=0.0,1.0"], seeds=2, baselines=True)
    assert [r["arm"] for r in rows] == ["centralized", "no-privacy", "configured"] * 2
    assert all(r["final_accuracy"] is not None and r["final_accuracy_std"] is not None for r in rows)
    dp_row = [r for r in rows if r["arm"] == "configured" and r["fl.sigma"] == 1.0][0]
    assert dp_row["epsilon"] is not None and dp_row["epsilon"] > 0
    table = format_table(rows)
    assert "+-" in table and "epsilon" in table and len(table.splitlines()) == len(rows) + 2
    json.dumps(rows)


def test_auc_matches_the_rank_definition():
    s = torch.tensor([0.1, 0.4, 0.35, 0.8])
    y = torch.tensor([0, 0, 1, 1])
    assert roc_auc(s, y == 1) == pytest.approx(0.75)
    assert roc_auc(torch.zeros(4), y == 1) == pytest.approx(0.5)          # all tied
    logits = torch.tensor([[2.0, 0.1, 0.0], [0.0, 3.0, 0.1], [0.2, 0.1, 1.0], [1.5, 0.0, 0.3]])
    assert auc_from_logits(logits, torch.tensor([0, 1, 2, 0])) == pytest.approx(1.0)


def test_key_agreement_masks_cancel_and_late_dropouts_are_repaired():
    K, D = 6, 40
    ka = KeyAgreement(K, seed=11)
    t = ka.table()
    assert (t == t.T).all() and len({int(v) for v in t[np.triu_indices(K, 1)]}) == K * (K - 1) // 2
    assert all(ka.shared_secret(i, j) == ka.shared_secret(j, i) for i in range(K) for j in range(i))
    assert KeyAgreement(K, seed=12).table()[0, 1] != t[0, 1]
    g = torch.Generator().manual_seed(0)
    vals = torch.randn(K, D, generator=g).double().numpy() * 0.1
    msgs = mask_messages(vals, seed=5, rnd=3, participants=range(K), pair_keys=t)
    with np.errstate(over="ignore"):
        assert np.allclose(decode_fixed(msgs.sum(0, dtype=np.uint64)), vals.sum(0), atol=1e-8)
        # the masks really depend on the pair secrets (different table -> different messages, same sum)
        other = mask_messages(vals, seed=5, rnd=3, participants=range(K), pair_keys=KeyAgreement(K, 12).table())
        assert (other != msgs).any() and np.allclose(decode_fixed(other.sum(0, dtype=np.uint64)), vals.sum(0), atol=1e-8)
        # clients 1 and 4 vanish after everybody masked against them: the plain sum of the survivors is garbage ...
        surv, late = [0, 2, 3, 5], [1, 4]
        total = msgs[surv].sum(0, dtype=np.uint64)
        assert not np.allclose(decode_fixed(total), vals[surv].sum(0), atol=1e-3)
        # ... until the survivors reveal those pair secrets and the server strips the orphaned masks
        fixed = unmask_dropped(total, 5, 3, surv, late, K, pair_keys=t)
        assert np.allclose(decode_fixed(fixed), vals[surv].sum(0), atol=1e-8)
        assert all(ka.reveal(s, d) == int(t[s, d]) for s in surv for d in late)


def test_runner_late_dropouts_with_key_agreement_equal_the_plain_average():
    """The protocol path (mask against the trained set, drop, reveal, unmask) gives the survivors' plain weighted mean."""
    torch.manual_seed(0)
    K = 6
    clients = []
    for k in range(K):
        g = torch.Generator().manual_seed(50 + k)
        y = torch.randint(0, 2, (12 + k,), generator=g)
        clients.append((torch.rand(12 + k, 4, generator=g) * 0.4 + y[:, None] * 1.8, y))
    outs = {}
    for secagg in (False, True):
        torch.manual_seed(1)
        model = VQC(4, 2, 2)
        data = FederatedDataset.from_clients(clients, "cpu")
        tr = TorchLocalTrainer(model, data, None, epochs=1, batch_size=64, lr=0.1, optimizer="sgd", momentum=0.0,
                               n_wrap=model.n_angle_params)
        cfg = FLConfig(num_rounds=3, mode="delta", late_drop_prob=0.4, secagg=secagg, secagg_keys="dh", seed=9, quiet=True)
        runner = FederatedRunner(tr, cfg, None)
        for _ in range(3):
            runner.run_round()
        outs[secagg] = (runner.global_flat.clone(), runner.revealed_secrets)
    assert torch.allclose(outs[True][0], outs[False][0], atol=1e-6)
    assert outs[True][1] > 0 and outs[False][1] == 0


def test_dp_accounting_status_and_fixed_denominator():
    torch.manual_seed(0)
    clients = [(torch.rand(10, 4), torch.randint(0, 2, (10,))) for _ in range(8)]

    def runner(**kw):
        model = VQC(4, 1, 2)
        tr = TorchLocalTrainer(model, FederatedDataset.from_clients(clients, "cpu"), None, epochs=1, batch_size=16, lr=0.1,
                               optimizer="sgd", momentum=0.0, n_wrap=model.n_angle_params)
        return FederatedRunner(tr, FLConfig(num_rounds=4, mode="delta", clip=1.0, sigma=1.0, quiet=True, **kw), None)

    strict = runner(client_fraction=0.5, poisson_sampling=True)
    loose = runner(client_fraction=0.5, poisson_sampling=False)
    assert strict.epsilon_status() == "rdp-sampled-gaussian" and loose.epsilon_status() == "indicative"
    for _ in range(4):
        strict.run_round()
    # the accountant is fed the configured rate q, never the realised fraction, and z = sigma for uniform weights
    assert all(q == 0.5 and z == pytest.approx(1.0) for q, z, _n in strict.accountant.history)
    assert strict.accountant.epsilon(1e-5) > 0
    out = runner(client_fraction=1.0).run()
    assert out["epsilon_kind"] == "rdp-sampled-gaussian" and out["comm_mb_total"] > 0


@pytest.mark.parametrize("world", [2])
def test_cli_train_under_torchrun_matches_single_process(tmp_path, world):
    """`torchrun -m qfedx_b200 train` spreads the clients over the ranks and reproduces the single-process accuracies."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    args = ["train", "--preset", "iris4", "--set", "device=cpu", "--set", "fl.quiet=true", "--set", "fl.num_rounds=3",
            "--set", "train.local_epochs=1", "--set", "train.optimizer=sgd", "--set", "train.batch_size=64",
            "--set", "train.momentum=0.0"]
    single = subprocess.run([sys.executable, "-m", "qfedx_b200"] + args, capture_output=True, text=True, timeout=300,
                            env=env, cwd=ROOT)
    assert single.returncode == 0, single.stderr[-2000:]
    multi = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                            "--master-addr", "127.0.0.1", "--master-port", "29671", "-m", "qfedx_b200"] + args,
                           capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert multi.returncode == 0, multi.stdout[-2000:] + multi.stderr[-2000:]
    a = json.loads([l for l in single.stdout.splitlines() if l.startswith("{")][-1])
    b = json.loads([l for l in multi.stdout.splitlines() if l.startswith("{")][-1])
    assert b["world"] == world and a["world"] == 1
    assert a["accuracies"] == pytest.approx(b["accuracies"], abs=1e-6)


def test_cluster_tiers_follow_the_step_counts():
    """Per-client cluster sizes of the fused VQC trainer (pure planning, no GPU): the longest client gets the widest
    cluster, nobody gets a wider one than it needs to keep up, equal clients share one tier, and circuits narrower than a
    warp never go beyond the one-pass size."""
    from qfedx_b200.fl.trainers import plan_cluster_tiers
    steps = [175, 140, 100, 60, 40, 10, 5]
    tiers = plan_cluster_tiers(steps, 32, 4, 8)
    assert tiers[0] == 8 and tiers == sorted(tiers, reverse=True) and tiers[-1] == 1
    cost = lambda c: 2.7 + (32 + c - 1) // c
    for st, c in zip(steps, tiers):
        assert st * cost(c) <= steps[0] * cost(8) or c == 8
        if c > 1:
            assert st * cost(c // 2) > steps[0] * cost(8)            # the next narrower cluster would be too slow
    assert plan_cluster_tiers([50, 50, 50, 50], 32, 1, 32) == [1, 1, 1, 1]        # 4-qubit circuits: one CTA covers the batch
    assert set(plan_cluster_tiers([90, 90], 32, 4, 8)) == {8}
    assert plan_cluster_tiers([], 32, 4, 8) == []
