"""Variational quantum classifier as an ``nn.Module`` with an ordinary ``state_dict``.

Feature map RY(x_i) (x already scaled to angles, e.g. pi * minmax(features), the mapping of
/root/reference/src/QFed/qAngle.py:35-41) or an amplitude-encoded input state
(/root/reference/src/QFed/qAmplitude.py:25-41); hardware-efficient ansatz of L layers
[RX(theta) RZ(phi) per qubit + CNOT/CZ ring]; readout <Z_j> on qubit j for j < C;
``logit_j = a_j <Z_j> + b_j`` (/root/reference/ROADMAP.md:126-128).  C = 1 is the binary
classifier (BCE-with-logits), C > 1 uses softmax cross-entropy.

On CUDA the forward/backward run on the register-tiled sm_100a kernels through
``qfedx_b200.ops.statevector``; ``grad_method`` selects adjoint (K4a) or parameter shift (K4b).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import statevector as _sv


class VQC(nn.Module):
    def __init__(self, n_qubits: int = 4, n_layers: int = 2, num_classes: int = 2, basis: str = "ry",
                 entangler: str = "cnot", encoding: str = "angle", grad_method: str = "adjoint",
                 init_scale: float = 0.1, noise=None):
        super().__init__()
        self.n_qubits, self.n_layers = n_qubits, n_layers
        self.n_logits = 1 if num_classes <= 2 else num_classes
        if self.n_logits > n_qubits:
            raise ValueError(f"{num_classes} classes need at least {self.n_logits} readout qubits")
        self.num_classes = num_classes
        self.basis, self.entangler, self.encoding, self.grad_method = basis, entangler, encoding, grad_method
        # optional qfedx_b200.noise.NoiseConfig: readout-level noise (depolarizing / damping / confusion / shots) on <Z>
        self.noise = noise if (noise is not None and noise.enabled) else None
        self.theta = nn.Parameter(init_scale * torch.randn(n_layers, n_qubits, 2))
        self.readout_scale = nn.Parameter(torch.ones(self.n_logits))
        self.readout_bias = nn.Parameter(torch.zeros(self.n_logits))

    # flat layout used by the fused kernels: [theta | a | b]
    @property
    def n_angle_params(self) -> int:
        return self.theta.numel()

    def flat_params(self) -> torch.Tensor:
        return torch.cat([self.theta.detach().reshape(-1), self.readout_scale.detach(), self.readout_bias.detach()])

    def load_flat_params(self, flat: torch.Tensor) -> None:
        p, c = self.theta.numel(), self.n_logits
        with torch.no_grad():
            self.theta.copy_(flat[:p].reshape(self.theta.shape))
            self.readout_scale.copy_(flat[p:p + c])
            self.readout_bias.copy_(flat[p + c:p + 2 * c])

    def expvals(self, x: torch.Tensor) -> torch.Tensor:
        if self.encoding == "amplitude":
            return _sv.vqc_expvals(None, self.theta, None, self.basis, self.entangler, init_state=x,
                                   method=self.grad_method)
        return _sv.vqc_expvals(x, self.theta, None, self.basis, self.entangler, method=self.grad_method)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: [B, n] feature-map angles (or [B, 2^n] unit-norm amplitudes) -> logits [B, n_logits]."""
        E = self.expvals(x)[:, : self.n_logits]
        if self.noise is not None:
            from ..noise import apply_to_expvals
            E = apply_to_expvals(E, self.noise)      # affine channels keep the gradient, shots are straight-through
        return self.readout_scale * E + self.readout_bias

    def loss(self, logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.n_logits == 1:
            return F.binary_cross_entropy_with_logits(logits[:, 0], y.float())
        return F.cross_entropy(logits, y)

    @torch.no_grad()
    def inference_expvals(self, x: torch.Tensor, engine: str = "auto") -> torch.Tensor:
        """<Z_q> [B, n] without autograd, on the fastest engine for the shape: the register kernel K1 (n <= 10), the
        tcgen05 unitary GEMM K2 (deep circuits over large batches, also at 11-12 qubits where no register kernel exists: its cost does not grow with the
        depth) or the pass engine K5 (n > 12).  ``engine`` forces 'registers' | 'gemm' | 'passes'."""
        if not x.is_cuda or self.encoding != "angle":
            return self.expvals(x)
        from ..ops.unitary_gemm import MAX_QUBITS, UnitaryGemm, prefers_unitary_gemm
        n, L, B = self.n_qubits, self.n_layers, x.shape[0]
        if engine == "auto":
            if prefers_unitary_gemm(n, L, B):
                engine = "gemm"
            else:
                engine = "registers" if n <= _sv.MAX_REG_QUBITS else "passes"
        if engine == "gemm":
            key = (self.theta._version, self.theta.data_ptr(), self.entangler, self.basis)
            if getattr(self, "_ug_key", None) != key:     # the packed unitary is rebuilt only when theta changed
                self._ug = UnitaryGemm(self.theta.detach(), self.entangler, self.basis)
                self._ug_key = key
            return self._ug.expvals(x.float())
        if engine == "passes":
            from ..ops.large import LargeVQC
            if getattr(self, "_large", None) is None:
                self._large = LargeVQC(n, L, self.entangler)
            return self._large.expvals(x.float(), self.theta.detach().reshape(1, -1), None)
        return _sv.sv_forward(x, self.theta.detach(), None, self.basis, self.entangler)

    @torch.no_grad()
    def predict_logits(self, x: torch.Tensor, engine: str = "auto") -> torch.Tensor:
        E = self.inference_expvals(x, engine)[:, : self.n_logits]
        if self.noise is not None:
            from ..noise import apply_to_expvals
            E = apply_to_expvals(E, self.noise)
        return self.readout_scale * E + self.readout_bias

    @torch.no_grad()
    def predict(self, x: torch.Tensor, engine: str = "auto") -> torch.Tensor:
        z = self.predict_logits(x, engine)
        return (z[:, 0] > 0).long() if self.n_logits == 1 else z.argmax(1)
