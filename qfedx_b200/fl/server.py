"""Server loop: sampling -> local training -> fused aggregation -> evaluation -> accounting ->
checkpoint.  Counterpart of federated_learning (/root/reference/src/CFed/Classical_FL.py:104-157)
extended with the planned pieces of /root/reference/ROADMAP.md:29-62,90-93: client sampling,
delta aggregation with angle wrapping, DP clip + noise with an accountant, secure aggregation,
client-dropout tolerance, checkpoint every K rounds, per-round metric logging.

Unlike the reference's main(), which swallows every exception and prints it
(/root/reference/src/CFed/Classical_FL.py:216-218), failures propagate.
"""
from __future__ import annotations

import os
import time
from dataclasses import asdict, dataclass, field
from typing import Optional

import numpy as np
import torch

from ..ops.aggregate import fedavg_aggregate
from ..utils.checkpoint import latest_checkpoint, load_checkpoint, save_checkpoint
from ..utils.logging import RoundLogger
from ..utils.seed import rng_state, set_rng_state
from ..utils.tracing import PhaseTimer
from .accountant import RDPAccountant
from .sampler import inject_dropouts, sample_clients


@dataclass
class FLConfig:
    num_rounds: int = 30
    client_fraction: float = 1.0
    poisson_sampling: bool = False
    drop_prob: float = 0.0           # fault injection: each selected client drops with this probability
    mode: str = "weights"            # 'weights' (reference FedAvg) or 'delta'
    clip: float = 0.0                # L2 clip C (0 = off); needs mode='delta'
    sigma: float = 0.0               # noise multiplier
    secagg: bool = False
    server_lr: float = 1.0
    dp_delta: float = 1e-5
    seed: int = 42
    eval_every: int = 1
    checkpoint_every: int = 0
    checkpoint_dir: Optional[str] = None
    resume: bool = False
    log_path: Optional[str] = None
    tensorboard_dir: Optional[str] = None
    print_every: int = 5
    quiet: bool = False


class FederatedRunner:
    def __init__(self, trainer, cfg: FLConfig, test_data=None, aggregator=None, client_gids=None,
                 global_weights=None):
        """``client_gids``/``global_weights`` place this process in a multi-rank job: the trainer's
        local clients have these global ids and ``global_weights`` holds every client's sample
        count; sampling and dropouts are then decided over the global population (identically on
        every rank, they are pure functions of seed and round)."""
        self.trainer, self.cfg, self.test_data = trainer, cfg, test_data
        self.aggregator = aggregator  # optional multi-GPU aggregator with the same call signature
        self.client_gids = None if client_gids is None else [int(g) for g in client_gids]
        self.global_weights = global_weights
        self.global_flat = trainer.global_flat().clone()
        self.accountant = RDPAccountant()
        self.history = {"accuracy": [], "loss": [], "epsilon": [], "round_ms": []}
        self.round = 0
        self.timer = PhaseTimer()
        self.logger = RoundLogger(cfg.log_path, cfg.tensorboard_dir, cfg.print_every, cfg.quiet)
        self._weights = trainer.weights if global_weights is None else torch.as_tensor(global_weights, dtype=torch.float32)
        if cfg.resume and cfg.checkpoint_dir:
            ck = latest_checkpoint(cfg.checkpoint_dir)
            if ck:
                self.load(ck)

    # ---- persistence ---------------------------------------------------------------------
    def save(self, path=None) -> str:
        path = path or os.path.join(self.cfg.checkpoint_dir, f"round_{self.round:06d}.pt")
        self.trainer.load_flat(self.global_flat)
        return save_checkpoint(path, rnd=self.round, model_state=self.trainer.model.state_dict(),
                               flat_params=self.global_flat.detach().cpu(), accountant=self.accountant.state_dict(),
                               history=self.history, config=asdict(self.cfg), rng=rng_state())

    def load(self, path) -> None:
        ck = load_checkpoint(path)
        self.round = ck["round"]
        self.global_flat = ck["flat"].to(self.global_flat.device)
        self.trainer.load_flat(self.global_flat)
        if ck.get("accountant"):
            self.accountant.load_state_dict(ck["accountant"])
        self.history = ck.get("history") or self.history
        if ck.get("rng"):
            set_rng_state(ck["rng"])

    # ---- one round ---------------------------------------------------------------------------
    def evaluate(self):
        if self.test_data is None:
            return None, None
        X, y = self.test_data
        return self.trainer.evaluate(self.global_flat, X, y)

    def run_round(self) -> dict:
        cfg, rnd = self.cfg, self.round
        K = len(self._weights)
        selected = sample_clients(K, cfg.client_fraction, rnd, cfg.seed, cfg.poisson_sampling)
        survivors = inject_dropouts(selected, cfg.drop_prob, rnd, cfg.seed)
        n_wrap = self.trainer.n_wrap if cfg.mode == "delta" else 0
        common = dict(mode=cfg.mode, n_wrap=n_wrap, clip=cfg.clip, sigma=cfg.sigma, server_lr=cfg.server_lr,
                      seed=cfg.seed, rnd=rnd, secagg=cfg.secagg)
        if self.client_gids is None:
            with self.timer.phase("train_ms"):
                client_params, stats = self.trainer.train(self.global_flat, survivors, rnd)
            w = self._weights[torch.as_tensor(survivors, dtype=torch.long)]
            with self.timer.phase("aggregate_ms"):
                res = (self.aggregator or fedavg_aggregate)(client_params, self.global_flat, w, None, **common)
        else:
            alive = set(int(s) for s in survivors)
            rows = [i for i, g in enumerate(self.client_gids) if g in alive]
            with self.timer.phase("train_ms"):
                if rows:
                    client_params, stats = self.trainer.train(self.global_flat, rows, rnd)
                else:
                    client_params, stats = None, {}
            part = torch.zeros(K, dtype=torch.int32)
            part[torch.as_tensor(survivors, dtype=torch.long)] = 1
            w = self._weights[torch.as_tensor(survivors, dtype=torch.long)]
            with self.timer.phase("aggregate_ms"):
                res = self.aggregator(client_params, self.global_flat, self._weights, part,
                                      gids=[self.client_gids[i] for i in rows], **common)
        self.global_flat = res.new_global
        if cfg.mode == "delta" and cfg.clip > 0 and cfg.sigma > 0:
            # per-client noise sigma*C on each of m averaged updates: the aggregate carries noise of
            # std sigma*C*sqrt(sum w~^2) against a sensitivity of max_k w~_k * C
            wn = (w / w.sum()).double()
            eff_sigma = float(cfg.sigma * torch.sqrt((wn ** 2).sum()) / wn.max())
            self.accountant.step(min(1.0, len(selected) / K) if cfg.client_fraction < 1 else 1.0, eff_sigma)
        self.round += 1
        return {"selected": int(len(selected)), "survivors": int(len(survivors)), **{k: v for k, v in stats.items()
                                                                                    if isinstance(v, float)}}

    def run(self) -> dict:
        cfg = self.cfg
        if self.round == 0 and self.test_data is not None:
            loss0, acc0 = self.evaluate()
            self.history["accuracy"].append(acc0)
            self.history["loss"].append(loss0)
            self.logger.log(0, accuracy=acc0, loss=loss0)
        while self.round < cfg.num_rounds:
            t0 = time.perf_counter()
            info = self.run_round()
            metrics = dict(info)
            if self.test_data is not None and (self.round % cfg.eval_every == 0 or self.round == cfg.num_rounds):
                with self.timer.phase("eval_ms"):
                    loss, acc = self.evaluate()
                self.history["accuracy"].append(acc)
                self.history["loss"].append(loss)
                metrics.update(accuracy=acc, loss=loss)
            if cfg.clip > 0 and cfg.sigma > 0:
                eps = self.accountant.epsilon(cfg.dp_delta)
                self.history["epsilon"].append(eps)
                metrics["epsilon"] = eps
            ms = (time.perf_counter() - t0) * 1e3
            self.history["round_ms"].append(ms)
            metrics["round_ms"] = ms
            metrics["comm_mb"] = 2.0 * info["survivors"] * self.global_flat.numel() * 4 / 1e6
            metrics.update(self.timer.summary_ms())   # device ms per phase (train / aggregate / eval)
            self.logger.log(self.round, **metrics)
            if cfg.checkpoint_every and cfg.checkpoint_dir and self.round % cfg.checkpoint_every == 0:
                self.save()
        self.trainer.load_flat(self.global_flat)
        self.logger.close()
        return {"model": self.trainer.model, "accuracies": self.history["accuracy"], "history": self.history,
                "global_flat": self.global_flat}
