"""Phase breakdown of the VQC local-training kernel from its in-kernel cycle counters (config 3 of BASELINE.json).

usage: python scripts/profile_vqc_train.py [cluster_size [n_clients]]   (on a GPU box; uniform cluster size, default 4;
n_clients keeps only the largest clients)
Prints microseconds per optimizer step and phase for the client with the most steps and the mean over clients."""
import ctypes
import importlib.util
import sys

import torch

sys.path.insert(0, ".")
from qfedx_b200.fl import FederatedDataset
from qfedx_b200.fl.trainers import VQCFusedTrainer
from qfedx_b200.models import VQC
from qfedx_b200.ops import _lib

NAMES = ["set-up", "forward+<Z>", "loss+sinks", "backward", "cluster wait", "grad reduce", "cluster bar 2", "optimizer",
         "diag tables", "-", "-", "-"]


def main():
    cs = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    keep = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    spec = importlib.util.spec_from_file_location("bench", "bench.py")
    bench = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    spec.loader.exec_module(bench)
    sys.argv = argv
    cfg = bench.VQC8
    clients, _ = bench.vqc_data(cfg)
    if keep:
        clients = sorted(clients, key=lambda c: -len(c[0]))[:keep]
    dev = torch.device("cuda")
    data = FederatedDataset.from_clients([(c[0].float().contiguous(), c[1].long().contiguous()) for c in clients], dev)
    model = VQC(cfg["n_qubits"], cfg["n_layers"], cfg["classes"]).to(dev)
    tr = VQCFusedTrainer(model, data, cfg["epochs"], cfg["batch"], cfg["lr"], "adam", cluster_size=cs)
    K = len(clients)
    g0 = tr.global_flat()
    for r in range(2):
        tr.train(g0, list(range(K)), r)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); tr.train(g0, list(range(K)), 2); e.record(); torch.cuda.synchronize()
    steps = torch.tensor([cfg["epochs"] * ((len(c[0]) + cfg["batch"] - 1) // cfg["batch"]) for c in clients], dtype=torch.float64)
    print(f"cluster size {tr.cluster_size}: round {s.elapsed_time(e):.3f} ms, longest client {int(steps.max())} steps "
          f"-> {s.elapsed_time(e) * 1e3 / float(steps.max()):.1f} us/step")
    prof = torch.zeros(K, 12, dtype=torch.int64, device=dev)
    _lib.call("qfx_vqc_train_profile", ctypes.c_void_p(prof.data_ptr()), launches=0)
    tr.train(g0, list(range(K)), 3)
    torch.cuda.synchronize()
    _lib.call("qfx_vqc_train_profile", ctypes.c_void_p(0), launches=0)
    us = prof.double().cpu() / 1965.0 / steps[:, None]          # cycles -> us per step at 1965 MHz
    big = int(steps.argmax())
    print(f"{'phase':14s} {'longest':>8s} {'mean':>8s}   (us per optimizer step)")
    for i, n in enumerate(NAMES):
        if us[:, i].max() > 0:
            print(f"{n:14s} {us[big, i]:8.2f} {us[:, i].mean():8.2f}")
    print(f"{'sum':14s} {us[big].sum():8.2f} {us.sum(1).mean():8.2f}")


if __name__ == "__main__":
    main()
