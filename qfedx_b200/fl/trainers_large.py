"""Trainers for VQCs whose state does not fit in registers: the pass engine (K5, one GPU per state,
BASELINE.json config 4) and the sharded state (K6, one state over 2/4/8 GPUs, config 5).

Gradients come from the parameter-shift rule (/root/reference/ROADMAP.md:130-135) - the 2P shifted
circuits of a minibatch run as one batch of states - or from SPSA (/root/reference/ROADMAP.md:38), which
needs two circuit evaluations per step whatever the parameter count.  The readout
(logit_j = a_j <Z_j> + b_j, BCE / CE) and the optimizer are ordinary torch code on tiny tensors.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from ..models.vqc import VQC
from ..ops.large import LargeVQC, ansatz_gates, ring_readout_masks
from .dataset import FederatedDataset


class _ExpvalBackend:
    n_out = None   # qubits read out (None: all); the trainer sets it to the model's logit count

    def expvals(self, angles, theta):  # [S,n], [P] -> [S,n_out]
        raise NotImplementedError

    def shift_grad(self, angles, theta, cot):  # -> [P]
        raise NotImplementedError


class PassBackend(_ExpvalBackend):
    def __init__(self, n, L, entangler="cnot", max_states=256):
        self.vqc, self.max_states = LargeVQC(n, L, entangler), max_states

    def expvals(self, angles, theta):
        return self.vqc.expvals(angles, theta.reshape(1, -1), self.n_out)

    def shift_grad(self, angles, theta, cot):
        return self.vqc.param_shift_grad(angles, theta, cot, self.max_states)

    def adjoint_grad(self, angles, theta, cot):
        return self.vqc.adjoint_grad(angles, theta, cot[:, : self.n_out] if self.n_out else cot)


class ShardedBackend(_ExpvalBackend):
    """One circuit at a time, its state sharded over every rank of the job."""

    def __init__(self, n, L, world, rank, device, entangler="cnot", loopback=False):
        from ..parallel.sharded import ShardedState, plan_sharded
        self.n, self.L = n, L
        # the last ring is folded into the parity readout: its cross-shard CNOTs would cost another swap
        self.plan = plan_sharded(n, int(math.log2(world)), ansatz_gates(n, L, entangler, final_ring=False))
        self.masks = ring_readout_masks(n, entangler if L > 0 else "none")
        self.state = ShardedState(n, world, rank, device, loopback=loopback)

    def _one(self, ang, theta):
        self.state.reset_zero_state()
        self.state.run(self.plan, theta, ang.reshape(1, -1))
        return self.state.expvals_parity(self.plan, self.masks[:self.n_out or self.n])

    def expvals(self, angles, theta):
        return torch.stack([self._one(angles[s], theta) for s in range(angles.shape[0])])

    def shift_grad(self, angles, theta, cot):
        g = torch.zeros_like(theta)
        for p in range(theta.numel()):
            tp, tm = theta.clone(), theta.clone()
            tp[p] += math.pi / 2
            tm[p] -= math.pi / 2
            g[p] = 0.5 * ((self.expvals(angles, tp) - self.expvals(angles, tm)) * cot).sum()
        return g


class LargeVQCTrainer:
    def __init__(self, model: VQC, data: FederatedDataset, backend: _ExpvalBackend, epochs: int = 1,
                 batch_size: int = 4, lr: float = 0.05, grad_method: str = "param_shift", spsa_c: float = 0.1,
                 generator=None):
        self.model, self.data, self.backend = model, data, backend
        self.epochs, self.batch_size, self.lr = epochs, batch_size, lr
        self.grad_method, self.spsa_c, self.generator = grad_method, spsa_c, generator
        self.n_wrap = model.n_angle_params
        self.launches = 0
        backend.n_out = model.n_logits

    @property
    def dim(self):
        return self.model.n_angle_params + 2 * self.model.n_logits

    @property
    def weights(self):
        return torch.as_tensor(self.data.sizes, dtype=torch.float32)

    def global_flat(self):
        return self.model.flat_params().to(self.data.device)

    def load_flat(self, flat):
        self.model.load_flat_params(flat)

    def _loss_and_cot(self, E, a, b, y):
        C = self.model.n_logits
        Ec = E[:, :C].detach().requires_grad_(True)
        a_, b_ = a.detach().requires_grad_(True), b.detach().requires_grad_(True)
        z = a_ * Ec + b_
        loss = F.binary_cross_entropy_with_logits(z[:, 0], y.float()) if C == 1 else F.cross_entropy(z, y)
        gE, ga, gb = torch.autograd.grad(loss, [Ec, a_, b_])
        cot = torch.zeros_like(E)
        cot[:, :C] = gE
        return loss.detach(), cot, ga, gb

    def _train_one(self, flat, k):
        P, C = self.model.n_angle_params, self.model.n_logits
        w = flat.clone().float()
        m, v, t = torch.zeros_like(w), torch.zeros_like(w), 0
        X, y = self.data.client(k)
        last = 0.0
        for _ in range(self.epochs):
            order = torch.randperm(len(X), device=X.device, generator=self.generator)
            for lo in range(0, len(X), self.batch_size):
                idx = order[lo:lo + self.batch_size]
                xb, yb = X[idx].float(), y[idx]
                theta, a, b = w[:P], w[P:P + C], w[P + C:]
                E = self.backend.expvals(xb, theta)
                loss, cot, ga, gb = self._loss_and_cot(E, a, b, yb)
                if self.grad_method == "spsa":
                    delta = (torch.randint(0, 2, (P,), device=w.device, generator=self.generator).float() * 2 - 1)
                    lp = self._loss_and_cot(self.backend.expvals(xb, theta + self.spsa_c * delta), a, b, yb)[0]
                    lm = self._loss_and_cot(self.backend.expvals(xb, theta - self.spsa_c * delta), a, b, yb)[0]
                    gth = (lp - lm) / (2 * self.spsa_c) * delta
                elif self.grad_method == "adjoint":
                    if not hasattr(self.backend, "adjoint_grad"):
                        raise ValueError("adjoint differentiation is implemented for the pass engine (PassBackend)")
                    gth = self.backend.adjoint_grad(xb, theta, cot)
                else:
                    gth = self.backend.shift_grad(xb, theta, cot)
                g = torch.cat([gth, ga, gb])
                t += 1
                m = 0.9 * m + 0.1 * g
                v = 0.999 * v + 0.001 * g * g
                w = w - self.lr * (m / (1 - 0.9 ** t)) / (torch.sqrt(v / (1 - 0.999 ** t)) + 1e-8)
                last = float(loss)
        return w, last

    def _train_lockstep(self, flat, selected):
        """All selected clients of this rank advance together (pass engine + adjoint gradients): at step j every client
        that still has a minibatch contributes its samples to ONE batch of states - sample row s of client k sits at batch
        index s * K + k, so the kernels pick client k's parameter row with `index % K` - and one forward sweep, one readout
        and one backward sweep serve all of them; Adam runs vectorised over the clients.  Same arithmetic per client as the
        one-client-at-a-time loop (sample orders are drawn in that loop's order)."""
        P, C, dev = self.model.n_angle_params, self.model.n_logits, flat.device
        sel = [int(k) for k in selected]
        data = [self.data.client(k) for k in sel]
        K, bs = len(sel), self.batch_size
        orders = [[torch.randperm(len(X), device=X.device, generator=self.generator) for _ in range(self.epochs)] for X, _ in data]
        W = flat.float()[None].repeat(K, 1).contiguous()
        m, v = torch.zeros_like(W), torch.zeros_like(W)
        t = torch.zeros(K, device=dev)
        last = torch.zeros(K, device=dev)
        n_k = torch.tensor([len(X) for X, _ in data], device=dev)
        n_feat = data[0][0].shape[1]
        for e in range(self.epochs):
            for j in range(int(math.ceil(int(n_k.max()) / bs))):
                act = [i for i in range(K) if j * bs < len(data[i][0])]
                A = len(act)
                xb = torch.zeros(bs, A, n_feat, device=dev)
                yb = torch.zeros(bs, A, dtype=torch.long, device=dev)
                valid = torch.zeros(bs, A, device=dev)
                for c, i in enumerate(act):
                    idx = orders[i][e][j * bs:(j + 1) * bs]
                    s_ = len(idx)
                    xb[:s_, c], yb[:s_, c], valid[:s_, c] = data[i][0][idx].float(), data[i][1][idx], 1.0
                    if s_ < bs:                                 # pad with a real sample: zero cotangent, no contribution
                        xb[s_:, c] = data[i][0][idx[0]].float()
                rows = torch.as_tensor(act, device=dev)
                Wa = W[rows]
                theta, a, b = Wa[:, :P].contiguous(), Wa[:, P:P + C], Wa[:, P + C:]
                E = self.backend.vqc.expvals(xb.reshape(bs * A, n_feat), theta, C).reshape(bs, A, C)
                Ed = E.detach().requires_grad_(True)
                a_, b_ = a.detach().requires_grad_(True), b.detach().requires_grad_(True)
                z = a_[None] * Ed + b_[None]
                if C == 1:
                    per = F.binary_cross_entropy_with_logits(z[..., 0], yb.float(), reduction="none")
                else:
                    per = F.cross_entropy(z.reshape(bs * A, C), yb.reshape(-1), reduction="none").reshape(bs, A)
                loss_k = (per * valid).sum(0) / valid.sum(0)                 # every client's own minibatch mean
                gE, ga, gb = torch.autograd.grad(loss_k.sum(), [Ed, a_, b_])
                gth = self.backend.vqc.adjoint_grad(xb.reshape(bs * A, n_feat), theta, gE.reshape(bs * A, C).contiguous())
                g = torch.cat([gth, ga, gb], dim=1)
                t[rows] += 1
                m[rows] = 0.9 * m[rows] + 0.1 * g
                v[rows] = 0.999 * v[rows] + 0.001 * g * g
                tt = t[rows][:, None]
                W[rows] = Wa - self.lr * (m[rows] / (1 - 0.9 ** tt)) / (torch.sqrt(v[rows] / (1 - 0.999 ** tt)) + 1e-8)
                last[rows] = loss_k.detach()
        return W, {"train_loss": float(last.mean())}

    def train(self, global_flat, selected, rnd):
        if self.grad_method == "adjoint" and isinstance(self.backend, PassBackend) and global_flat.is_cuda and len(selected) > 1:
            return self._train_lockstep(global_flat, selected)
        rows, losses = [], []
        for k in selected:
            w, l = self._train_one(global_flat, int(k))
            rows.append(w)
            losses.append(l)
        return torch.stack(rows), {"train_loss": float(np.mean(losses)) if losses else 0.0}

    @torch.no_grad()
    def evaluate(self, global_flat, X, y, batch_size: int = 64):
        P, C = self.model.n_angle_params, self.model.n_logits
        theta, a, b = global_flat[:P], global_flat[P:P + C], global_flat[P + C:]
        from ..ops.unitary_gemm import UnitaryGemm, prefers_unitary_gemm
        if X.is_cuda and isinstance(self.backend, PassBackend) and prefers_unitary_gemm(self.model.n_qubits,
                                                                                          self.model.n_layers, len(X)):
            # 11-12 qubits, a deep circuit and a test set of hundreds of samples: one tcgen05 GEMM with the round's unitary (K2)
            ug = UnitaryGemm(theta.reshape(1, self.model.n_layers, self.model.n_qubits, 2), self.model.entangler)
            z = a * ug.expvals(X.float())[:, :C] + b
            loss = F.binary_cross_entropy_with_logits(z[:, 0], y.float()) if C == 1 else F.cross_entropy(z, y)
            pred = (z[:, 0] > 0).long() if C == 1 else z.argmax(1)
            return float(loss), float((pred == y).float().mean())
        tot, correct = 0.0, 0
        for lo in range(0, len(X), batch_size):
            xb, yb = X[lo:lo + batch_size].float(), y[lo:lo + batch_size]
            z = a * self.backend.expvals(xb, theta)[:, :C] + b
            tot += float(F.binary_cross_entropy_with_logits(z[:, 0], yb.float()) if C == 1 else F.cross_entropy(z, yb)) * len(xb)
            pred = (z[:, 0] > 0).long() if C == 1 else z.argmax(1)
            correct += int((pred == yb).sum())
        return tot / len(X), correct / len(X)
