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


def run_experiment(cfg: ExperimentConfig) -> dict:
    from .data import PCAReducer, angle_features, images_to_angles, preprocess_mnist
    from .fl import FederatedDataset, FederatedRunner, TorchLocalTrainer, VQCFusedTrainer
    from .models import VQC, TinyCNN
    set_seeds(cfg.seed)
    dev = _device(cfg.device)
    d = cfg.data
    prep = preprocess_mnist(d.raw_folder, d.processed_folder, tuple(d.digits), d.val_split, d.num_clients,
                            d.partition_type, d.alpha, remap_labels=d.remap_labels,
                            synthetic_fallback=d.synthetic_fallback, synthetic_scale=d.synthetic_scale,
                            verbose=not cfg.fl.quiet)
    if prep is None:
        raise SystemExit(f"no raw data under {d.raw_folder} and synthetic_fallback is off")
    n_classes = len(d.digits)
    if cfg.model.family == "tinycnn":
        data = FederatedDataset.from_clients(prep.clients, dev)
        model = TinyCNN(n_classes).to(dev)
        if dev.type == "cuda" and cfg.model.engine != "torch":
            from .fl.trainers_cnn import TinyCNNBatchedTrainer
            trainer = TinyCNNBatchedTrainer(model, data, cfg.train.local_epochs, cfg.train.batch_size, cfg.train.lr,
                                            cfg.train.momentum)
        else:
            trainer = TorchLocalTrainer(model, data, torch.nn.CrossEntropyLoss(), cfg.train.local_epochs,
                                        cfg.train.batch_size, cfg.train.lr, "sgd", cfg.train.momentum)
        test = (prep.test[0].to(dev), prep.test[1].to(dev))
    else:
        n = cfg.model.n_qubits
        if d.reducer == "pca":
            pca = PCAReducer(n).fit(prep.train[0].to(dev))
            feats = lambda X: angle_features(pca.transform(X.to(dev)), n)
        else:
            feats = lambda X: images_to_angles(X.to(dev), n)
        clients = [(feats(X), y.to(dev)) for X, y in prep.clients]
        data = FederatedDataset.from_clients(clients, dev)
        model = VQC(n, cfg.model.n_layers, n_classes, entangler=cfg.model.entangler,
                    grad_method=cfg.model.grad_method).to(dev)
        engine = cfg.model.engine
        if engine == "auto":
            engine = "registers" if n <= 10 else "passes"
        if engine == "registers" or dev.type != "cuda":
            if cfg.train.optimizer == "spsa":
                trainer = TorchLocalTrainer(model, data, None, cfg.train.local_epochs, cfg.train.batch_size,
                                            cfg.train.lr, "adam", n_wrap=model.n_angle_params, grad_estimator="spsa")
            else:
                trainer = VQCFusedTrainer(model, data, cfg.train.local_epochs, cfg.train.batch_size, cfg.train.lr,
                                          cfg.train.optimizer, cfg.train.momentum, grad_method=cfg.model.grad_method)
        else:
            from .fl.trainers_large import LargeVQCTrainer, PassBackend, ShardedBackend
            if engine == "sharded":
                import torch.distributed as dist
                world = dist.get_world_size() if dist.is_initialized() else 1
                rank = dist.get_rank() if dist.is_initialized() else 0
                if world == 1:   # single process: emulate the ranks in one device's memory
                    backend = ShardedBackend(n, cfg.model.n_layers, 2, 0, dev, cfg.model.entangler, loopback=True)
                else:
                    backend = ShardedBackend(n, cfg.model.n_layers, world, rank, dev, cfg.model.entangler)
            else:
                backend = PassBackend(n, cfg.model.n_layers, cfg.model.entangler)
            method = "spsa" if cfg.train.optimizer == "spsa" else "param_shift"
            trainer = LargeVQCTrainer(model, data, backend, cfg.train.local_epochs, cfg.train.batch_size,
                                      cfg.train.lr, method)
        test = (feats(prep.test[0]), prep.test[1].to(dev))
    out = FederatedRunner(trainer, cfg.fl, test).run()
    return {"name": cfg.name, "accuracies": out["accuracies"], "final_accuracy": out["accuracies"][-1] if out["accuracies"] else None,
            "epsilon": out["history"]["epsilon"][-1] if out["history"]["epsilon"] else None,
            "round_ms": out["history"]["round_ms"]}


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
    sub.add_parser("encode-demo")
    sub.add_parser("build")
    b = sub.add_parser("bench")
    b.add_argument("rest", nargs=argparse.REMAINDER)
    args = ap.parse_args(argv)
    if args.cmd == "train":
        cfg = load_config(args.config, args.overrides)
        if args.print_config:
            print(dump_config(cfg))
        print(json.dumps(run_experiment(cfg)))
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
