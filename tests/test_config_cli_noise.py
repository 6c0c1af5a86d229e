import json
import math

import numpy as np
import pytest
import torch

from qfedx_b200.cli import encode_demo, run_experiment
from qfedx_b200.config import dump_config, load_config
from qfedx_b200.noise import NoiseConfig, apply_to_expvals, trajectory_expvals
from qfedx_b200.ops import ref


def test_presets_load_and_override():
    for name in ("vqc4", "cfed", "vqc8", "vqc20", "vqc30"):
        cfg = load_config(name, ["fl.num_rounds=2", "train.lr=0.2"])
        assert cfg.name == name and cfg.fl.num_rounds == 2 and cfg.train.lr == 0.2
    assert load_config("vqc8").data.partition_type == "non_iid"
    assert "n_qubits" in dump_config(load_config())
    with pytest.raises(KeyError):
        load_config(None, ["model.no_such_key=1"])


def test_cli_train_small_vqc_cpu(tmp_path):
    cfg = load_config("vqc4", ["device=cpu", "data.synthetic_scale=0.02", f"data.raw_folder={tmp_path}",
                               f"data.processed_folder={tmp_path / 'p'}", "fl.num_rounds=2", "fl.quiet=true",
                               "train.local_epochs=1", "data.reducer=block"])
    out = run_experiment(cfg)
    assert len(out["accuracies"]) == 3 and json.dumps(out)


def test_encode_demo_matches_reference_recipe(capsys):
    out = encode_demo()
    assert out["amplitudes"].shape == (16,) and float(out["amplitudes"].norm()) == pytest.approx(1.0)
    ang = out["angles"]
    assert float(ang.min()) == 0.0 and float(ang.max()) == pytest.approx(math.pi)
    assert "RY(" in capsys.readouterr().out


def test_readout_noise_closed_forms():
    E = torch.tensor([[0.8, -0.4]])
    assert torch.allclose(apply_to_expvals(E, NoiseConfig("depolarizing", p=0.1)), 0.9 * E)
    out = apply_to_expvals(E, NoiseConfig(readout_p01=0.05, readout_p10=0.1))
    p0 = 0.5 * (1 + E)
    assert torch.allclose(out, 2 * (p0 * 0.95 + (1 - p0) * 0.1) - 1)
    g = torch.Generator().manual_seed(0)
    shots = apply_to_expvals(E.repeat(4000, 1), NoiseConfig(shots=100), g)
    assert shots.mean(0)[0] == pytest.approx(0.8, abs=0.01)
    assert shots.std(0)[0] == pytest.approx(math.sqrt((1 - 0.64) / 100), rel=0.1)


def test_noise_degrades_signal():
    g = torch.Generator().manual_seed(1)
    ang = torch.rand(6, 3, generator=g) * math.pi
    theta = 0.3 * torch.randn(1, 2, 3, 2, generator=g)
    ideal = ref.vqc_expvals_ref(ang, theta)
    noisy = trajectory_expvals(ang, theta, NoiseConfig("depolarizing", p=0.15), n_traj=200, generator=g)
    damp = trajectory_expvals(ang, theta, NoiseConfig("amplitude", gamma=0.2), n_traj=100, generator=g)
    assert noisy.abs().mean() < ideal.abs().mean()             # depolarizing shrinks <Z> towards 0
    assert damp.mean() > ideal.mean()                          # amplitude damping pulls towards |0>, <Z> = +1
    zero = trajectory_expvals(ang, theta, NoiseConfig("depolarizing", p=0.0), n_traj=2, generator=g)
    assert torch.allclose(zero, ideal, atol=1e-5)
