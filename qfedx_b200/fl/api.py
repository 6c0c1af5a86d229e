"""User-facing entry points.

``federated_learning`` keeps the call shape of the reference's function of the same name
(/root/reference/src/CFed/Classical_FL.py:104-110: client_data, test_data, num_rounds,
local_epochs, learning_rate, batch_size, num_classes -> {'model', 'accuracies'}) and adds the
model family, the privacy options and the device.  ``quantum_federated_learning`` is the VQC
pipeline of /root/reference/ROADMAP.md:29-43 with the reference's image -> 4x4 -> pooled-angles
feature path (/root/reference/src/QFed/testEncoder.py:94,124-125).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from ..data.features import PCAReducer, angle_features, images_to_angles
from ..models import VQC, TinyCNN
from .dataset import FederatedDataset
from .server import FederatedRunner, FLConfig
from .trainers import TorchLocalTrainer, VQCFusedTrainer

TensorPair = Tuple[torch.Tensor, torch.Tensor]


def _device(device):
    if device is None:
        return torch.device("cuda" if torch.cuda.is_available() else "cpu")
    return torch.device(device)


def federated_learning(client_data: Sequence[TensorPair], test_data: Optional[TensorPair] = None,
                       num_rounds: int = 30, local_epochs: int = 5, learning_rate: float = 0.01,
                       batch_size: int = 32, num_classes: int = 3, device=None, fl_config: FLConfig | None = None,
                       engine: str = "auto", **fl_kwargs) -> dict:
    """Classical FedAvg with TinyCNN.  ``engine``: 'auto' picks the client-batched CUDA path on a
    GPU and the eager torch loop on CPU; 'torch' forces the eager loop."""
    dev = _device(device)
    cfg = fl_config or FLConfig(num_rounds=num_rounds, **fl_kwargs)
    data = FederatedDataset.from_clients(client_data, dev)
    model = TinyCNN(num_classes).to(dev)
    test = None if test_data is None else (test_data[0].to(dev), test_data[1].to(dev).long())
    trainer = None
    if engine != "torch" and dev.type == "cuda":
        from .trainers_cnn import TinyCNNBatchedTrainer
        trainer = TinyCNNBatchedTrainer(model, data, epochs=local_epochs, batch_size=batch_size, lr=learning_rate,
                                        seed=cfg.seed)
    if trainer is None:
        trainer = TorchLocalTrainer(model, data, torch.nn.CrossEntropyLoss(), local_epochs, batch_size,
                                    learning_rate, "sgd", 0.9)
    return FederatedRunner(trainer, cfg, test).run()


def prepare_quantum_features(X: torch.Tensor, n_qubits: int, reducer: str = "block",
                             pca: PCAReducer | None = None) -> torch.Tensor:
    """Images [N,1,28,28] -> feature-map angles [N, n_qubits] in [0, pi]."""
    if reducer == "block":
        return images_to_angles(X, n_qubits)
    if reducer == "pca":
        return angle_features(pca.transform(X), n_qubits)
    raise ValueError(reducer)


def quantum_federated_learning(client_data: Sequence[TensorPair], test_data: Optional[TensorPair] = None,
                               n_qubits: int = 4, n_layers: int = 2, num_classes: int = 2, num_rounds: int = 30,
                               local_epochs: int = 5, learning_rate: float = 0.05, batch_size: int = 32,
                               optimizer: str = "adam", reducer: str = "block", grad_method: str = "adjoint",
                               device=None, fl_config: FLConfig | None = None, entangler: str = "cnot",
                               noise=None, **fl_kwargs) -> dict:
    """VQC FedAvg.  Inputs are images (reduced to angles here) or already [N, n_qubits] angles.
    ``noise``: an optional qfedx_b200.noise.NoiseConfig applied to the readout during training and evaluation."""
    dev = _device(device)
    if fl_config is None:
        fl_kwargs.setdefault("mode", "delta")
        cfg = FLConfig(num_rounds=num_rounds, **fl_kwargs)
    else:
        cfg = fl_config
    pca = None
    if reducer == "pca" and client_data[0][0].dim() > 2:
        allX = torch.cat([c[0] for c in client_data]).to(dev)
        pca = PCAReducer(n_qubits).fit(allX)

    def feats(X):
        X = X.to(dev)
        return X.float() if X.dim() == 2 and X.shape[1] == n_qubits else prepare_quantum_features(X, n_qubits, reducer, pca)

    clients = [(feats(X), y.to(dev).long()) for X, y in client_data]
    data = FederatedDataset.from_clients(clients, dev)
    model = VQC(n_qubits, n_layers, num_classes, entangler=entangler, grad_method=grad_method, noise=noise).to(dev)
    test = None if test_data is None else (feats(test_data[0]), test_data[1].to(dev).long())
    trainer = VQCFusedTrainer(model, data, local_epochs, batch_size, learning_rate, optimizer,
                              grad_method=grad_method, seed=cfg.seed, noise=noise)
    out = FederatedRunner(trainer, cfg, test).run()
    out["pca"] = pca
    return out
