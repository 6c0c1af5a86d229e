"""Phase breakdown of the persistent TinyCNN round kernel (K8p) from its in-kernel cycle counters.

usage: python scripts/profile_cnn_round.py [clients] [ctas_per_client]   (on a GPU box)
Prints microseconds per step and phase: mean over the CTAs and the slowest CTA."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from qfedx_b200.data import synthetic_mnist
from qfedx_b200.fl import FederatedDataset
from qfedx_b200.fl.trainers_cnn import TinyCNNBatchedTrainer
from qfedx_b200.models import TinyCNN
from qfedx_b200.ops import _lib

NAMES = ["img load", "conv1", "fwd mma", "fwd epi", "bar1", "fc1 fwd", "bar2", "head", "fc1 bwd", "bar3", "planes",
         "dgrad mma", "dg epi+c1wg", "wgrad wait", "wgrad epi", "wload", "bar4", "update", "bar5", "head:load+h",
         "head:logits+CE", "head:dh", "-", "-"]


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    G = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    per = 4190
    X, y = synthetic_mnist(K * per, 3, seed=0)
    data = FederatedDataset.from_clients([(X[k * per:(k + 1) * per], y[k * per:(k + 1) * per]) for k in range(K)], "cuda")
    tr = TinyCNNBatchedTrainer(TinyCNN(3).cuda(), data, epochs=5, batch_size=32, lr=0.01, ctas_per_client=G)
    g0 = tr.global_flat()
    for _ in range(2):
        tr.train(g0, list(range(K)), 0)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); tr.train(g0, list(range(K)), 1); e.record(); torch.cuda.synchronize()
    steps = 5 * ((per + 31) // 32)
    print(f"clients {K}: round {s.elapsed_time(e):.2f} ms, {s.elapsed_time(e) * 1e3 / steps:.1f} us/step")
    geff = G if G else min(32, torch.cuda.get_device_properties(0).multi_processor_count // K)
    prof = torch.zeros(K, geff, 24, dtype=torch.int64, device="cuda")
    _lib.call("qfx_tinycnn_round_profile", ctypes.c_void_p(prof.data_ptr()), launches=0)
    tr.train(g0, list(range(K)), 2)
    torch.cuda.synchronize()
    _lib.call("qfx_tinycnn_round_profile", ctypes.c_void_p(0), launches=0)
    us = prof.double().cpu() / 1965.0 / steps          # cycles -> us per step at 1965 MHz
    flat = us.view(-1, 24)
    print(f"{'phase':14s} {'mean':>8s} {'max':>8s}   (us per step, G = {geff} CTAs per client)")
    for i, n in enumerate(NAMES):
        if flat[:, i].max() > 0:
            print(f"{n:14s} {flat[:, i].mean():8.2f} {flat[:, i].max():8.2f}")
    print(f"{'sum':14s} {flat.sum(1).mean():8.2f} {flat.sum(1).max():8.2f}")


if __name__ == "__main__":
    main()
