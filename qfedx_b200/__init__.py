"""qfedx_b200 - Blackwell-native quantum federated learning.

PyTorch is the framework layer (models, optimizers, state_dict checkpoints, orchestration);
the VQC statevector engine, the local-training loop, the federated aggregation and the
classical client-batched training run on hand-written sm_100a CUDA kernels
(qfedx_b200/csrc, C ABI in include/qfedx_b200.h).
"""
__version__ = "0.1.0"

from .utils.seed import set_seeds  # noqa: F401
