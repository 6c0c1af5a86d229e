"""Local-training engines.  A trainer turns (global flat parameters, selected clients, round) into
the matrix of client parameters [K_sel, D] that the aggregation consumes.

* ``TorchLocalTrainer``  - any ``nn.Module``; sequential clients, fresh optimizer per round,
  full passes over shuffled minibatches: the semantics of the reference's client_update
  (/root/reference/src/CFed/Classical_FL.py:40-64).  CPU path, numerics reference, and baseline.
* ``VQCFusedTrainer``    - all selected clients in ONE launch of csrc/vqc_train.cu on CUDA
  (one thread-block cluster per client); on CPU it defers to ``TorchLocalTrainer``.
"""
from __future__ import annotations

import copy
from typing import Callable, Optional

import numpy as np
import torch
import torch.nn as nn

from ..models.tinycnn import flatten_state, unflatten_state
from ..models.vqc import VQC
from ..ops import _lib
from ..ops.statevector import BASIS, ENTANGLER
from .dataset import FederatedDataset
from .spsa import spsa_gradient

OPTIMIZERS = {"sgd": 0, "adam": 1}


def _check_labels(y: torch.Tensor, n_classes: int, what: str) -> None:
    """The kernels index logits by label: anything outside [0, n_classes) is rejected up front (torch's losses raise
    for the same input)."""
    if y.numel():
        lo, hi = int(y.min()), int(y.max())
        if lo < 0 or hi >= max(2, n_classes):
            raise _lib.QfxError(f"{what}: labels must lie in [0, {max(2, n_classes)}), found [{lo}, {hi}]")


def make_optimizer(params, name: str, lr: float, momentum: float = 0.9, betas=(0.9, 0.999), eps: float = 1e-8):
    if name == "sgd":
        return torch.optim.SGD(params, lr=lr, momentum=momentum)
    if name == "adam":
        return torch.optim.Adam(params, lr=lr, betas=betas, eps=eps)
    raise ValueError(f"unknown optimizer {name!r}")


class TorchLocalTrainer:
    def __init__(self, model: nn.Module, data: FederatedDataset, loss_fn: Optional[Callable] = None,
                 epochs: int = 5, batch_size: int = 32, lr: float = 0.01, optimizer: str = "sgd",
                 momentum: float = 0.9, n_wrap: int = 0, grad_estimator: str = "backprop", spsa_c: float = 0.1,
                 generator=None):
        self.model, self.data = model, data
        self.loss_fn = loss_fn or (model.loss if hasattr(model, "loss") else nn.CrossEntropyLoss())
        self.epochs, self.batch_size, self.lr = epochs, batch_size, lr
        self.optimizer, self.momentum = optimizer, momentum
        self.n_wrap = n_wrap
        self.grad_estimator, self.spsa_c = grad_estimator, spsa_c
        self.generator = generator
        self._template = copy.deepcopy(model.state_dict())
        self.launches = 0

    @property
    def dim(self) -> int:
        return sum(v.numel() for v in self._template.values())

    @property
    def weights(self) -> torch.Tensor:
        return torch.as_tensor(self.data.sizes, dtype=torch.float32)

    def global_flat(self) -> torch.Tensor:
        return flatten_state(self.model.state_dict())

    def load_flat(self, flat: torch.Tensor) -> None:
        self.model.load_state_dict(unflatten_state(flat, self._template))

    def _train_one(self, flat: torch.Tensor, k: int) -> tuple:
        self.load_flat(flat)
        self.model.train()
        Xk, yk = self.data.client(k)
        opt = make_optimizer(self.model.parameters(), self.optimizer, self.lr, self.momentum)
        last_loss, n = 0.0, len(Xk)
        for _ in range(self.epochs):
            order = torch.randperm(n, device=Xk.device, generator=self.generator)
            tot = 0.0
            for lo in range(0, n, self.batch_size):
                idx = order[lo:lo + self.batch_size]
                xb, yb = Xk[idx], yk[idx]
                opt.zero_grad(set_to_none=True)
                if self.grad_estimator == "spsa":
                    loss = spsa_gradient(self.model, lambda: self.loss_fn(self.model(xb), yb), self.spsa_c)
                else:
                    loss = self.loss_fn(self.model(xb), yb)
                    loss.backward()
                opt.step()
                tot += float(loss.detach()) * len(idx)
            last_loss = tot / max(n, 1)
        return flatten_state(self.model.state_dict()).clone(), last_loss

    def train(self, global_flat: torch.Tensor, selected, rnd: int):
        rows, losses = [], []
        for k in selected:
            f, l = self._train_one(global_flat, int(k))
            rows.append(f)
            losses.append(l)
        self.load_flat(global_flat)
        return torch.stack(rows), {"train_loss": float(np.mean(losses)) if losses else 0.0}

    @torch.no_grad()
    def evaluate(self, global_flat: torch.Tensor, X: torch.Tensor, y: torch.Tensor, batch_size: int = 256):
        self.load_flat(global_flat)
        self.model.eval()
        tot_loss, correct = 0.0, 0
        for lo in range(0, len(X), batch_size):
            xb, yb = X[lo:lo + batch_size], y[lo:lo + batch_size]
            z = self.model(xb)
            tot_loss += float(self.loss_fn(z, yb)) * len(xb)
            pred = (z[:, 0] > 0).long() if z.shape[1] == 1 else z.argmax(1)
            correct += int((pred == yb).sum())
        return tot_loss / len(X), correct / len(X)


def plan_cluster_tiers(steps, batch_size: int, one_pass_size: int, slots_per_cta: int):
    """Cluster size (CTAs) per client of a round of the fused VQC trainer, from the clients' optimizer-step counts.

    A round lasts as long as its slowest client, and a client's steps are sequential, so the client with the most steps
    gets the widest cluster (the fewest circuits per SM and step) and every other client the NARROWEST cluster that still
    finishes within that client's time - small clients then take one or two SMs instead of four, and the job stays within
    one wave of CTAs.  Cost model of a step, measured with the kernel's phase timers (scripts/profile_vqc_train.py,
    8 qubits x 6 layers): a fixed part worth about 2.7 circuits plus one unit per circuit a CTA simulates.

    ``one_pass_size``: the cluster that covers a minibatch in one pass; ``slots_per_cta``: circuits a CTA simulates at
    once (8 when a circuit is a whole warp, n >= 6 qubits - only then is a cluster of twice the one-pass size useful:
    halving the warps per SM still shortens the step, measured 24.1 -> 20.6 us; narrower circuits share warps and gain
    nothing from more CTAs)."""
    top = min(8, one_pass_size * (2 if slots_per_cta == 8 else 1))
    sizes = [c for c in (1, 2, 4, 8) if c <= top]
    cost = {c: 2.7 + (batch_size + c - 1) // c for c in sizes}
    target = max(steps) * cost[sizes[-1]] if len(steps) else 0.0
    return [next(c for c in sizes if st * cost[c] <= target or c == sizes[-1]) for st in steps]


class VQCFusedTrainer:
    """All selected clients train in one kernel launch (CUDA); see csrc/vqc_train.cu."""

    def __init__(self, model: VQC, data: FederatedDataset, epochs: int = 5, batch_size: int = 32,
                 lr: float = 0.05, optimizer: str = "adam", momentum: float = 0.9, betas=(0.9, 0.999),
                 eps: float = 1e-8, grad_method: str = "adjoint", cluster_size: int = 0, shuffle: bool = True,
                 generator=None, lane_mode: str = "latency", shuffle_mode: str = "feistel", seed: int = 0, noise=None):
        if model.encoding != "angle":
            raise ValueError("the fused trainer takes angle-encoded features")
        _check_labels(data.y, model.num_classes, "client data")
        self.model, self.data = model, data
        self.epochs, self.batch_size, self.lr = epochs, batch_size, lr
        self.optimizer, self.momentum, self.betas, self.eps = optimizer, momentum, betas, eps
        self.grad_method, self.shuffle, self.generator = grad_method, shuffle, generator
        # 'feistel': the kernel derives each epoch's sample order from a keyed permutation (no array, no sort);
        # 'array': one device sort per round produces explicit permutations
        self.shuffle_mode, self.seed = shuffle_mode, seed
        self.n_wrap = model.n_angle_params
        # readout-level noise (qfedx_b200.noise.NoiseConfig): applied inside the kernels' <Z> epilogue on CUDA, through
        # the model's forward on the CPU fallback
        from ..noise import readout_affine
        self.noise = noise if (noise is not None and noise.enabled) else getattr(model, "noise", None)
        if self.noise is not None and self.noise.kind != "none" and self.noise.kind not in ("depolarizing", "amplitude"):
            raise ValueError(f"unknown noise kind {self.noise.kind!r}")
        self._noise_ab = readout_affine(self.noise) if self.noise is not None else (1.0, 0.0)
        self._shots = int(self.noise.shots) if self.noise is not None else 0
        if self.noise is not None:
            model.noise = self.noise
        self.cuda = data.device.type == "cuda"
        self.launches = 0
        if self.cuda:
            self.angles = data.X.contiguous().float()
            self.labels = data.y.to(torch.int32).contiguous()
            n = model.n_qubits
            self.lane_mode = 1 if lane_mode == "latency" else 0
            if self.lane_mode == 1:
                lane_bits = 0 if n <= 1 else min(n - 1, 5)
            else:
                lane_bits = 0 if n <= 4 else (n - 4 if n <= 9 else 5)
            slots = 256 >> lane_bits
            # cluster_size 0: per-client cluster sizes, chosen every round from the selected clients' step counts
            # (``_plan_tiers``); ``self.cluster_size`` is then the single-pass size used when all clients get the same
            self.tiered = cluster_size <= 0
            if cluster_size <= 0:  # enough CTAs to cover one minibatch in a single pass, at most 8
                cluster_size = 1
                while cluster_size < 8 and slots * cluster_size < batch_size:
                    cluster_size *= 2
            self.cluster_size = cluster_size
            self._slots = slots
            self._tier_streams = {}
            self._plans = {}
            self._out = None
        else:
            self._fallback = TorchLocalTrainer(model, data, None, epochs, batch_size, lr, optimizer, momentum,
                                               n_wrap=self.n_wrap, generator=generator)

    @property
    def dim(self) -> int:
        return self.model.n_angle_params + 2 * self.model.n_logits

    @property
    def weights(self) -> torch.Tensor:
        return torch.as_tensor(self.data.sizes, dtype=torch.float32)

    def global_flat(self) -> torch.Tensor:
        return self.model.flat_params().to(self.data.device)

    def load_flat(self, flat):
        self.model.load_flat_params(flat)

    def _plan_tiers(self, selected):
        bs = self.batch_size
        steps = [self.epochs * ((int(self.data.sizes[int(k)]) + bs - 1) // bs) for k in selected]
        return plan_cluster_tiers(steps, bs, self.cluster_size, self._slots)

    def train(self, global_flat: torch.Tensor, selected, rnd: int):
        if not self.cuda:
            return self._fallback.train(global_flat, selected, rnd)
        m = self.model
        selected = [int(k) for k in np.asarray(selected).reshape(-1)]
        K = len(selected)
        dev = self.angles.device
        # plan of a selection (cluster sizes, launch order, device index tensors): kept for the selections seen last - with
        # full participation every round reuses one entry and starts without host-to-device copies
        key = tuple(selected)
        plan = self._plans.get(key)
        if plan is None:
            cs_of = self._plan_tiers(selected) if self.tiered else [self.cluster_size] * K
            # clients ordered by tier (widest clusters first): each tier is one launch writing its slice of the outputs
            order = sorted(range(K), key=lambda i: -cs_of[i])
            sel = torch.as_tensor(np.asarray([selected[i] for i in order]), dtype=torch.int32, device=dev)
            inv = None if order == list(range(K)) else torch.as_tensor(np.argsort(np.asarray(order)), dtype=torch.long, device=dev)
            if len(self._plans) >= 16:
                self._plans.pop(next(iter(self._plans)))
            plan = self._plans[key] = (cs_of, order, sel, inv)
        cs_of, order, sel, inv = plan
        out = torch.empty(K, self.dim, device=dev, dtype=torch.float32)
        stats = torch.empty(K, 2, device=dev, dtype=torch.float32)
        from .prng import shuffle_key
        use_array = self.shuffle and self.shuffle_mode == "array"
        perm = self.data.epoch_permutations(self.epochs, self.generator) if use_array else None
        skey = shuffle_key(self.seed, rnd) if (self.shuffle and not use_array) else 0
        beta1 = self.momentum if self.optimizer == "sgd" else self.betas[0]
        g = global_flat.contiguous().float()
        from .prng import shuffle_key as _key
        tiers, lo = [], 0
        while lo < K:
            hi = lo
            while hi < K and cs_of[order[hi]] == cs_of[order[lo]]:
                hi += 1
            tiers.append((cs_of[order[lo]], lo, hi))
            lo = hi

        def launch(cs, lo, hi):
            _lib.call("qfx_vqc_local_train_noisy", _lib.c_int(m.n_qubits), _lib.c_int(m.n_layers), _lib.c_int(BASIS[m.basis]),
                      _lib.c_int(ENTANGLER[m.entangler]), _lib.c_int(m.n_logits), _lib.ptr(self.angles),
                      _lib.ptr(self.labels), _lib.ptr(self.data.offsets), _lib.ptr(perm),
                      _lib.c_int(self.angles.shape[0]), _lib.ptr(g), _lib.ptr(sel[lo:hi]), _lib.c_int(hi - lo),
                      _lib.ptr(out[lo:hi]), _lib.ptr(stats[lo:hi]), _lib.c_int(self.epochs), _lib.c_int(self.batch_size),
                      _lib.c_int(OPTIMIZERS[self.optimizer]), _lib.c_float(self.lr), _lib.c_float(beta1),
                      _lib.c_float(self.betas[1]), _lib.c_float(self.eps),
                      _lib.c_int(0 if self.grad_method == "adjoint" else 1), _lib.c_int(cs),
                      _lib.c_int(self.lane_mode), _lib.c_u64(skey), _lib.c_float(self._noise_ab[0]),
                      _lib.c_float(self._noise_ab[1]), _lib.c_int(self._shots), _lib.c_u64(_key(self.seed ^ 0x5EED, rnd)),
                      _lib.stream_ptr())
            self.launches += 1

        if len(tiers) == 1:
            launch(*tiers[0])
        else:
            # one launch per tier, side by side on their own streams (the tiers are independent; together they fill the GPU)
            main = torch.cuda.current_stream()
            fork = torch.cuda.Event()
            fork.record(main)
            for cs, lo, hi in tiers:
                st = self._tier_streams.get(cs)
                if st is None:
                    st = self._tier_streams[cs] = torch.cuda.Stream(device=dev)
                st.wait_event(fork)
                with torch.cuda.stream(st):
                    launch(cs, lo, hi)
                    done = torch.cuda.Event()
                    done.record(st)
                main.wait_event(done)
        if inv is not None:                # back to the caller's client order
            out, stats = out.index_select(0, inv), stats.index_select(0, inv)
        self.last_cluster_sizes = cs_of
        self._last_stats = stats
        return out, {"train_stats": stats}

    def evaluate(self, global_flat: torch.Tensor, X: torch.Tensor, y: torch.Tensor, batch_size: int = 0):
        _check_labels(y, self.model.num_classes, "evaluation data")
        if not self.cuda:
            return self._fallback.evaluate(global_flat, X, y)
        m = self.model
        out2 = torch.empty(2, device=X.device, dtype=torch.float32)
        self._eval_calls = getattr(self, "_eval_calls", 0) + 1
        _lib.call("qfx_vqc_eval_noisy", _lib.c_int(m.n_qubits), _lib.c_int(m.n_layers), _lib.c_int(BASIS[m.basis]),
                  _lib.c_int(ENTANGLER[m.entangler]), _lib.c_int(m.n_logits), _lib.ptr(X.contiguous().float()),
                  _lib.ptr(y.to(torch.int32).contiguous()), _lib.ptr(global_flat.contiguous().float()),
                  _lib.c_int(X.shape[0]), _lib.ptr(out2), _lib.c_float(self._noise_ab[0]),
                  _lib.c_float(self._noise_ab[1]), _lib.c_int(self._shots),
                  _lib.c_u64((self.seed * 0x9E3779B97F4A7C15 + self._eval_calls) & (2 ** 64 - 1)), _lib.stream_ptr())
        self.launches += 1
        res = out2.cpu()  # the round's single host sync
        return float(res[0]) / X.shape[0], float(res[1]) / X.shape[0]
