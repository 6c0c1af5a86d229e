"""Command line: ``python -m qfedx_b200 {train,preprocess,encode-demo,bench,build}``.

``train`` runs a preset or YAML config end to end (data -> partition -> federated training -> report);
``preprocess`` is the reference's Preprocess.py main (/root/reference/src/CFed/Preprocess.py:231-247);
``encode-demo`` is its testEncoder.py (/root/reference/src/QFed/testEncoder.py:58-129) without qiskit:
it prints the 4x4 image, the amplitude statevector and the angle-encoding rotation angles.
"""
from __future__ import annotations

import argparse
import json
import math
import random
import sys

import numpy as np
import torch

from .config import ExperimentConfig, dump_config, load_config
from .utils.seed import set_seeds


def _device(name: str) -> torch.device:
    if name == "auto":
        return torch.device("cuda" if torch.cuda.is_available() else "cpu")
    return torch.device(name)


def _dist_context():
    """(rank, world, local_rank) of a torchrun job, initialising the process group on first use; (0, 1, 0) otherwise."""
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1, 0
    import torch.distributed as dist
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group("gloo")
    return rank, world, local


def _prepare(cfg: ExperimentConfig):
    from .data import prepare_iris, preprocess_mnist
    d = cfg.data
    if d.dataset == "iris":
        return prepare_iris(tuple(d.digits), d.num_clients, d.partition_type, d.alpha, seed=cfg.seed)
    if d.dataset != "mnist":
        raise SystemExit(f"unknown dataset {d.dataset!r} (mnist | iris)")
    prep = preprocess_mnist(d.raw_folder, d.processed_folder, tuple(d.digits), d.val_split, d.num_clients,
                            d.partition_type, d.alpha, remap_labels=d.remap_labels,
                            synthetic_fallback=d.synthetic_fallback, synthetic_scale=d.synthetic_scale,
                            verbose=not cfg.fl.quiet)
    if prep is None:
        raise SystemExit(f"no raw data under {d.raw_folder} and synthetic_fallback is off")
    return prep


def run_experiment(cfg: ExperimentConfig) -> dict:
    """One experiment end to end.  Under ``torchrun`` (WORLD_SIZE > 1) the clients are spread over the ranks
    (parallel/runtime.py: local_clients), every rank trains its own clients and the round's aggregation runs through
    the cross-GPU fused aggregator (or the all_reduce baseline on CPU / gloo); every rank returns the same result.
    ``train.centralized`` pools all client data into one "client" - the centralized baseline of
    /root/reference/ROADMAP.md:4 - with everything else unchanged."""
    from .data import PCAReducer, angle_features, images_to_angles
    from .fl import FederatedDataset, FederatedRunner, TorchLocalTrainer, VQCFusedTrainer
    from .models import VQC, TinyCNN
    from .parallel.runtime import local_clients, make_aggregator
    from .utils.metrics import auc_from_logits
    rank, world, local = _dist_context()
    if rank != 0:      # one narrator per job: the other ranks neither print per-round lines nor write the round log
        cfg.fl.quiet, cfg.fl.log_path, cfg.fl.tensorboard_dir = True, None, None
    set_seeds(cfg.seed)
    dev = _device(cfg.device)
    if dev.type == "cuda" and world > 1:
        dev = torch.device("cuda", local)
    d = cfg.data
    prep = _prepare(cfg)
    n_classes = len(d.digits)
    noise = cfg.noise if cfg.noise.enabled else None
    clients = list(prep.clients)
    if cfg.train.centralized:
        clients = [(torch.cat([c[0] for c in clients]), torch.cat([c[1] for c in clients]))]
    K = len(clients)
    all_sizes = torch.tensor([len(c[0]) for c in clients], dtype=torch.float32)
    mine = local_clients(rank, K, world) if world > 1 else list(range(K))
    if world > 1 and not mine:
        raise SystemExit(f"rank {rank} has no clients: {K} clients cannot be spread over {world} ranks")
    local_data = [clients[k] for k in mine]
    if cfg.model.family == "tinycnn":
        if noise is not None:
            raise SystemExit("noise models apply to the VQC family")
        data = FederatedDataset.from_clients(local_data, dev)
        model = TinyCNN(n_classes).to(dev)
        if dev.type == "cuda" and cfg.model.engine != "torch":
            from .fl.trainers_cnn import TinyCNNBatchedTrainer
            trainer = TinyCNNBatchedTrainer(model, data, cfg.train.local_epochs, cfg.train.batch_size, cfg.train.lr,
                                            cfg.train.momentum, seed=cfg.seed,
                                            engine="graph" if cfg.model.engine == "graph" else "persistent")
        else:
            trainer = TorchLocalTrainer(model, data, torch.nn.CrossEntropyLoss(), cfg.train.local_epochs,
                                        cfg.train.batch_size, cfg.train.lr, "sgd", cfg.train.momentum)
        feats = lambda X: X.to(dev)
    else:
        n = cfg.model.n_qubits
        if d.dataset == "iris":
            if n != 4:
                raise SystemExit("the Iris task has 4 features: set model.n_qubits=4")
            feats = lambda X: X.to(dev)                       # prepare_iris already produced angles
        elif d.reducer == "pca":
            pca = PCAReducer(n).fit(prep.train[0].to(dev))
            feats = lambda X: angle_features(pca.transform(X.to(dev)), n)
        else:
            feats = lambda X: images_to_angles(X.to(dev), n)
        data = FederatedDataset.from_clients([(feats(X), y.to(dev)) for X, y in local_data], dev)
        model = VQC(n, cfg.model.n_layers, n_classes, entangler=cfg.model.entangler,
                    grad_method=cfg.model.grad_method, noise=noise).to(dev)
        engine = cfg.model.engine
        if engine == "auto":
            engine = "registers" if n <= 10 else "passes"
        if engine == "registers" or dev.type != "cuda":
            if cfg.train.optimizer == "spsa":
                trainer = TorchLocalTrainer(model, data, None, cfg.train.local_epochs, cfg.train.batch_size,
                                            cfg.train.lr, "adam", n_wrap=model.n_angle_params, grad_estimator="spsa")
            else:
                trainer = VQCFusedTrainer(model, data, cfg.train.local_epochs, cfg.train.batch_size, cfg.train.lr,
                                          cfg.train.optimizer, cfg.train.momentum, grad_method=cfg.model.grad_method,
                                          seed=cfg.seed, noise=noise)
        else:
            if noise is not None:
                raise SystemExit("readout noise is implemented for the register engine (n_qubits <= 10)")
            from .fl.trainers_large import LargeVQCTrainer, PassBackend, ShardedBackend
            if engine == "sharded":
                if world == 1:   # single process: emulate the ranks in one device's memory
                    backend = ShardedBackend(n, cfg.model.n_layers, 2, 0, dev, cfg.model.entangler, loopback=True)
                else:            # ONE state sharded over the ranks: every rank walks all clients in lockstep
                    backend = ShardedBackend(n, cfg.model.n_layers, world, rank, dev, cfg.model.entangler)
                    data = FederatedDataset.from_clients([(feats(X), y.to(dev)) for X, y in clients], dev)
                    mine = list(range(K))
            else:
                backend = PassBackend(n, cfg.model.n_layers, cfg.model.entangler)
            method = "spsa" if cfg.train.optimizer == "spsa" else cfg.model.grad_method
            trainer = LargeVQCTrainer(model, data, backend, cfg.train.local_epochs, cfg.train.batch_size,
                                      cfg.train.lr, method)
    test = (feats(prep.test[0]), prep.test[1].to(dev))
    sharded_job = cfg.model.family == "vqc" and cfg.model.engine == "sharded" and world > 1
    if world > 1 and not sharded_job:
        agg = make_aggregator(rank, world, mine, all_sizes, trainer.dim, dev, kind="fused" if dev.type == "cuda" else "nccl")
        runner = FederatedRunner(trainer, cfg.fl, test, aggregator=agg, client_gids=mine, global_weights=all_sizes)
    else:
        runner = FederatedRunner(trainer, cfg.fl, test)
    out = runner.run()
    hist = out["history"]
    # AUC of the final global model on the test split (one forward through the model API)
    auc = None
    try:
        model.eval()
        with torch.no_grad():
            auc = auc_from_logits(model(test[0]).float(), test[1])
    except Exception:   # engines without an autograd forward for this size (e.g. 30 sharded qubits)
        auc = None
    total_ms = float(sum(hist["round_ms"]))
    res = {"name": cfg.name, "accuracies": out["accuracies"],
           "final_accuracy": out["accuracies"][-1] if out["accuracies"] else None,
           "final_loss": hist["loss"][-1] if hist["loss"] else None, "final_auc": auc,
           "epsilon": hist["epsilon"][-1] if hist["epsilon"] else None, "epsilon_kind": out.get("epsilon_kind"),
           "round_ms": hist["round_ms"], "wall_clock_s": total_ms / 1e3,
           "comm_mb_per_round": hist["comm_mb"][-1] if hist["comm_mb"] else None, "comm_mb_total": out.get("comm_mb_total"),
           "gpu_hours": (total_ms / 3.6e6) * world if dev.type == "cuda" else 0.0,
           "world": world, "clients": K, "centralized": bool(cfg.train.centralized)}
    return res


def encode_demo(seed: int = 42) -> dict:
    from .data import (amplitude_normalize, angle_features, downsample_images, pool_features, synthetic_mnist)
    set_seeds(seed)
    X, y = synthetic_mnist(1, 3, seed=seed)
    small = downsample_images(X[0, 0].double(), (4, 4))
    flat16 = small.reshape(-1)
    amp = amplitude_normalize(flat16[None])[0]
    feats4 = pool_features(flat16[None], 4)[0]
    ang = angle_features(feats4[None])[0]
    np.set_printoptions(precision=4, suppress=True)
    print(f"label {int(y[0])}\n4x4 block means:\n{small.numpy()}")
    print(f"amplitude statevector (first 8 of 16 amplitudes): {amp[:8].numpy()}")
    print(f"angle encoding: RY angles on 4 qubits = {ang.numpy()}  (min feature -> 0, max -> pi)")
    for q, a in enumerate(ang.tolist()):
        print(f"  q{q}: --RY({a:.4f})--")
    return {"amplitudes": amp, "angles": ang}


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="qfedx_b200", description=__doc__)
    sub = ap.add_subparsers(dest="cmd", required=True)
    t = sub.add_parser("train")
    t.add_argument("--preset", "--config", dest="config", default=None)
    t.add_argument("--set", dest="overrides", action="append", default=[], metavar="key=value")
    t.add_argument("--print-config", action="store_true")
    p = sub.add_parser("preprocess")
    p.add_argument("--preset", "--config", dest="config", default=None)
    p.add_argument("--set", dest="overrides", action="append", default=[])
    p.add_argument("--plots", action="store_true")
    w = sub.add_parser("sweep", help="grid over config keys x seeds -> mean +- std table (ROADMAP.md:98-121)")
    w.add_argument("--preset", "--config", dest="config", default=None)
    w.add_argument("--set", dest="overrides", action="append", default=[], metavar="key=value")
    w.add_argument("--grid", action="append", default=[], metavar="key=v1,v2,...")
    w.add_argument("--seeds", type=int, default=3)
    w.add_argument("--out", default=None, help="write the rows as JSON here")
    w.add_argument("--baselines", action="store_true",
                   help="also run the centralized and the no-privacy federated arms of every grid point")
    sub.add_parser("encode-demo")
    sub.add_parser("build")
    b = sub.add_parser("bench")
    b.add_argument("rest", nargs=argparse.REMAINDER)
    args = ap.parse_args(argv)
    if args.cmd == "train":
        cfg = load_config(args.config, args.overrides)
        if args.print_config:
            print(dump_config(cfg))
        res = run_experiment(cfg)
        if res.get("world", 1) == 1 or int(__import__("os").environ.get("RANK", "0")) == 0:
            print(json.dumps(res))
        if res.get("world", 1) > 1:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
    elif args.cmd == "sweep":
        from .sweep import format_table, run_sweep
        rows = run_sweep(args.config, args.overrides, args.grid, args.seeds, baselines=args.baselines)
        print(format_table(rows))
        if args.out:
            from pathlib import Path
            Path(args.out).parent.mkdir(parents=True, exist_ok=True)
            Path(args.out).write_text(json.dumps(rows, indent=1))
    elif args.cmd == "preprocess":
        from .data import preprocess_mnist
        cfg = load_config(args.config, args.overrides)
        random.seed(cfg.seed); np.random.seed(cfg.seed)
        d = cfg.data
        preprocess_mnist(d.raw_folder, d.processed_folder, tuple(d.digits), d.val_split, d.num_clients,
                         d.partition_type, d.alpha, synthetic_fallback=d.synthetic_fallback,
                         synthetic_scale=d.synthetic_scale, plots=args.plots)
    elif args.cmd == "encode-demo":
        encode_demo()
    elif args.cmd == "build":
        from .build import build
        print(build())
    elif args.cmd == "bench":
        import runpy
        sys.argv = ["bench.py"] + args.rest
        runpy.run_path(str(__import__("pathlib").Path(__file__).resolve().parent.parent / "bench.py"), run_name="__main__")
    return 0


if __name__ == "__main__":
    sys.exit(main())
