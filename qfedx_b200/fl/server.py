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
    late_drop_prob: float = 0.0      # ... after it trained and its peers masked against it (secure-aggregation repair)
    mode: str = "weights"            # 'weights' (reference FedAvg) or 'delta'
    clip: float = 0.0                # L2 clip C (0 = off); needs mode='delta'
    sigma: float = 0.0               # noise multiplier
    secagg: bool = False
    secagg_keys: str = "job-seed"    # 'job-seed': pair masks from the shared job seed; 'dh': per-pair secrets from the
                                     # registration-time key agreement (fl/secagg.py: KeyAgreement)
    server_lr: float = 1.0
    dp_delta: float = 1e-5
    dp_fixed_denominator: bool = True  # DP-FedAvg estimator: sum_S w_k delta_k / (q W) instead of / sum_S w_k
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
        self.history = {"accuracy": [], "loss": [], "epsilon": [], "round_ms": [], "comm_mb": []}
        self.epsilon_kind = None
        self.key_agreement = None
        self.pair_keys = None
        if cfg.secagg and cfg.secagg_keys == "dh":
            from .secagg import KeyAgreement
            K_all = len(trainer.weights if global_weights is None else global_weights)
            self.key_agreement = KeyAgreement(K_all, cfg.seed)     # registration: every rank derives the same table
            self.pair_keys = self.key_agreement.table()
        elif cfg.secagg and cfg.secagg_keys != "job-seed":
            raise ValueError(f"secagg_keys must be 'job-seed' or 'dh', got {cfg.secagg_keys!r}")
        self.revealed_secrets = 0
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
        dp_on = cfg.mode == "delta" and cfg.clip > 0 and cfg.sigma > 0
        # under DP accounting an empty Poisson sample is a legitimate outcome (forcing a client in would break the
        # sampling assumption of the accountant): the round then leaves the model untouched
        selected = sample_clients(K, cfg.client_fraction, rnd, cfg.seed, cfg.poisson_sampling,
                                  allow_empty=dp_on and cfg.poisson_sampling)
        survivors = inject_dropouts(selected, cfg.drop_prob, rnd, cfg.seed) if len(selected) else selected
        late = np.zeros(0, dtype=np.int64)
        if cfg.late_drop_prob > 0 and len(survivors) > 1:
            kept = inject_dropouts(survivors, cfg.late_drop_prob, rnd, cfg.seed ^ 0x4C415445)
            late = np.setdiff1d(survivors, kept)
            trained, survivors = survivors, kept
        else:
            trained = survivors
        if len(survivors) == 0:
            self.round += 1
            if dp_on:
                self.accountant.step(cfg.client_fraction, self._noise_multiplier())
            return {"selected": 0, "survivors": 0}
        n_wrap = self.trainer.n_wrap if cfg.mode == "delta" else 0
        server_lr = cfg.server_lr
        w_alive = self._weights[torch.as_tensor(survivors, dtype=torch.long)]
        if dp_on and cfg.dp_fixed_denominator:
            # the kernel divides by the survivors' weight; rescale to the fixed denominator q * W_total so that the
            # estimator (and with it the sensitivity max_k w_k C / (q W)) does not depend on the sampled set
            q = min(1.0, cfg.client_fraction)
            server_lr = cfg.server_lr * float(w_alive.sum()) / (q * float(self._weights.sum()))
        common = dict(mode=cfg.mode, n_wrap=n_wrap, clip=cfg.clip, sigma=cfg.sigma, server_lr=server_lr,
                      seed=cfg.seed, rnd=rnd, secagg=cfg.secagg)
        if cfg.secagg and self.pair_keys is not None:
            common["pair_keys"] = self.pair_keys
        if cfg.secagg and len(late):
            # survivors disclose the secrets they share with the late drop-outs; the server strips those masks
            self.revealed_secrets += len(survivors) * len(late)
        if self.client_gids is None:
            with self.timer.phase("train_ms"):
                client_params, stats = self.trainer.train(self.global_flat, trained, rnd)
            if len(late):
                keep_rows = torch.as_tensor(np.flatnonzero(np.isin(trained, survivors)), dtype=torch.long,
                                            device=client_params.device)
                client_params = client_params.index_select(0, keep_rows)
            w = w_alive
            with self.timer.phase("aggregate_ms"):
                if self.aggregator is None and cfg.secagg and len(late) and not client_params.is_cuda:
                    # CPU reference plays the protocol literally: rows indexed by global id, masks against trained set
                    full = torch.zeros(K, client_params.shape[1], dtype=client_params.dtype)
                    full[torch.as_tensor(survivors, dtype=torch.long)] = client_params
                    part = torch.zeros(K, dtype=torch.int32)
                    part[torch.as_tensor(survivors, dtype=torch.long)] = 1
                    res = fedavg_aggregate(full, self.global_flat, self._weights, part, late_dropped=list(late), **common)
                else:
                    res = (self.aggregator or fedavg_aggregate)(client_params, self.global_flat, w, None, **common)
        else:
            alive = set(int(s) for s in survivors)
            was_trained = set(int(s) for s in trained)
            rows = [i for i, g in enumerate(self.client_gids) if g in was_trained]
            with self.timer.phase("train_ms"):
                if rows:
                    client_params, stats = self.trainer.train(self.global_flat, rows, rnd)
                else:
                    client_params, stats = None, {}
            if len(late) and rows:
                keep = [j for j, i in enumerate(rows) if self.client_gids[i] in alive]
                rows = [rows[j] for j in keep]
                client_params = client_params[torch.as_tensor(keep, dtype=torch.long, device=client_params.device)] if keep else None
            part = torch.zeros(K, dtype=torch.int32)
            part[torch.as_tensor(survivors, dtype=torch.long)] = 1
            w = w_alive
            with self.timer.phase("aggregate_ms"):
                res = self.aggregator(client_params, self.global_flat, self._weights, part,
                                      gids=[self.client_gids[i] for i in rows], **common)
        self.global_flat = res.new_global
        if dp_on:
            self.accountant.step(self._sampling_rate(len(selected), K), self._noise_multiplier(w))
        self.round += 1
        return {"selected": int(len(selected)), "survivors": int(len(survivors)), **{k: v for k, v in stats.items()
                                                                                    if isinstance(v, float)}}

    # ---- privacy accounting ---------------------------------------------------------------------
    def _sampling_rate(self, n_selected: int, K: int) -> float:
        return 1.0 if self.cfg.client_fraction >= 1 else float(self.cfg.client_fraction)

    def _noise_multiplier(self, w=None) -> float:
        """Noise multiplier z = (std of the aggregate's noise) / (add/remove sensitivity of the aggregate).
        Every update carries N(0, sigma^2 C^2) scaled by its weight; with the fixed denominator D = q W the aggregate's
        noise std is sigma C sqrt(sum_S w_k^2) / D and the sensitivity max_k w_k C / D, so z >= sigma * min_k w_k /
        max_k w_k for ANY non-empty sample - the bound used here (uniform weights: z = sigma)."""
        cfg = self.cfg
        wa = self._weights.double()
        if cfg.dp_fixed_denominator:
            return float(cfg.sigma * wa.min() / wa.max())
        # survivors' normalisation: the denominator moves with the sample, add/remove sensitivity is up to 2 max w~ C
        wn = (w / w.sum()).double()
        return float(cfg.sigma * torch.sqrt((wn ** 2).sum()) / (2.0 * wn.max()))

    def epsilon_status(self) -> str:
        """'rdp-sampled-gaussian' when the run matches the accountant's assumptions (Poisson sampling at a fixed rate
        or full participation, fixed denominator, no fault injection), 'indicative' otherwise.  Either way the noise
        and mask streams derive from the job seed that all simulated parties share: a simulation of the mechanism,
        not a deployment of it (privacy)."""
        cfg = self.cfg
        ok = (cfg.client_fraction >= 1 or cfg.poisson_sampling) and cfg.dp_fixed_denominator and \
            cfg.drop_prob == 0 and cfg.late_drop_prob == 0
        return "rdp-sampled-gaussian" if ok else "indicative"

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
                metrics["epsilon_kind"] = self.epsilon_status()
            ms = (time.perf_counter() - t0) * 1e3
            self.history["round_ms"].append(ms)
            metrics["round_ms"] = ms
            # down-link to every selected client + up-link from every survivor, fp32 parameters
            metrics["comm_mb"] = (info["selected"] + info["survivors"]) * self.global_flat.numel() * 4 / 1e6
            self.history["comm_mb"].append(metrics["comm_mb"])
            metrics["comm_mb_total"] = float(sum(self.history["comm_mb"]))
            metrics.update(self.timer.summary_ms())   # device ms per phase (train / aggregate / eval)
            self.logger.log(self.round, **metrics)
            if cfg.checkpoint_every and cfg.checkpoint_dir and self.round % cfg.checkpoint_every == 0:
                self.save()
        self.trainer.load_flat(self.global_flat)
        self.logger.close()
        return {"model": self.trainer.model, "accuracies": self.history["accuracy"], "history": self.history,
                "global_flat": self.global_flat, "comm_mb_total": float(sum(self.history["comm_mb"])),
                "epsilon_kind": self.epsilon_status() if (cfg.clip > 0 and cfg.sigma > 0) else None,
                "revealed_secrets": self.revealed_secrets}
