"""Runs each headline kernel a few times at (or near) its benchmark shape - the workload of the ncu captures.

usage: python scripts/profile_kernels.py [all|sv|gemm|agg|agg_loop|large|adjoint|swap|cnn|vqc]
Shapes are kept small enough that ncu's ~40 replays per launch stay in seconds; the persistent TinyCNN round kernel is
profiled on a short round (8 clients x 96 samples, one epoch = 24 steps)."""
import math
import sys

import torch

sys.path.insert(0, ".")
from qfedx_b200.ops import statevector as sv
from qfedx_b200.ops.aggregate import fedavg_aggregate
from qfedx_b200.ops.large import LargeVQC, ansatz_gates
from qfedx_b200.ops.unitary_gemm import UnitaryGemm
from qfedx_b200.parallel.sharded import ShardedState, plan_sharded

dev = torch.device("cuda")
torch.manual_seed(0)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "sv"):
    ang = torch.rand(1 << 15, 8, device=dev) * math.pi
    theta = torch.randn(1, 6, 8, 2, device=dev)
    cot = torch.randn(1 << 15, 8, device=dev)
    for _ in range(3):
        sv.sv_forward(ang, theta)
        sv.sv_backward(ang, theta, cot)
if which in ("all", "gemm"):
    ug = UnitaryGemm(torch.randn(1, 6, 8, 2, device=dev))
    ang = torch.rand(1 << 15, 8, device=dev) * math.pi
    for _ in range(3):
        ug.expvals(ang)                                   # <Z> only: no state store (the product's inference path)
    ug12 = UnitaryGemm(torch.randn(1, 4, 12, 2, device=dev))
    ang12 = torch.rand(4096, 12, device=dev) * math.pi
    for _ in range(2):
        ug12.expvals(ang12)
if which in ("all", "agg"):
    cp = torch.randn(8, 113859, device=dev)
    g = torch.randn(113859, device=dev)
    w = torch.rand(8, device=dev) + 1
    for _ in range(3):
        fedavg_aggregate(cp, g, w, mode="delta", clip=1.0, sigma=0.5, secagg=True)
    cp = torch.randn(4, 25_000_000, device=dev)
    g = torch.randn(25_000_000, device=dev)
    w = torch.rand(4, device=dev) + 1
    for _ in range(3):
        fedavg_aggregate(cp, g, w, mode="delta", clip=1.0)
    for _ in range(3):
        fedavg_aggregate(cp, g, w, mode="weights")
if which in ("all", "agg_loop"):
    sys.path.insert(0, "tests")
    from test_gpu_agg_loopback import loopback_aggregate
    K, D = 8, 4_000_000
    glob = torch.randn(D)
    cp_all = glob + 0.05 * torch.randn(K, D)
    for two_shot in (False, True):
        loopback_aggregate(4, cp_all, glob, torch.ones(K), torch.ones(K, dtype=torch.int32), two_shot, epochs=2,
                           mode="delta", n_wrap=0, clip=1.0, server_lr=1.0, seed=1, rnd=0, secagg=False)
if which in ("all", "large"):
    vqc = LargeVQC(20, 2)
    ang = torch.rand(16, 20, device=dev) * math.pi
    theta = torch.randn(1, vqc.n_params, device=dev)
    for _ in range(2):
        vqc.expvals(ang, theta)
if which in ("all", "adjoint"):
    vqc = LargeVQC(20, 2)
    ang = torch.rand(4, 20, device=dev) * math.pi
    theta = torch.randn(vqc.n_params, device=dev)
    cot = torch.randn(4, 1, device=dev)
    for _ in range(2):
        vqc.adjoint_grad(ang, theta, cot)
if which in ("all", "swap"):
    n, world = 27, 2          # 2 x 512 MiB shards in one device's memory (loopback)
    plan = plan_sharded(n, 1, ansatz_gates(n, 1))
    st = ShardedState(n, world, loopback=True)
    st.reset_zero_state()
    st.run(plan, torch.randn(2 * n, device=dev), torch.rand(1, n, device=dev))
if which in ("all", "cnn"):
    from qfedx_b200.data import synthetic_mnist
    from qfedx_b200.fl import FederatedDataset
    from qfedx_b200.fl.trainers_cnn import TinyCNNBatchedTrainer
    from qfedx_b200.models import TinyCNN
    K, per = 8, 96
    X, y = synthetic_mnist(K * per, 3, seed=0)
    data = FederatedDataset.from_clients([(X[k * per:(k + 1) * per], y[k * per:(k + 1) * per]) for k in range(K)], dev)
    tr = TinyCNNBatchedTrainer(TinyCNN(3).to(dev), data, epochs=1, batch_size=32, lr=0.01)
    for r in range(2):
        out, _ = tr.train(tr.global_flat(), list(range(K)), r)
    tr.evaluate(out[0], X[:512].to(dev), y[:512].to(dev))
if which in ("all", "vqc"):
    from qfedx_b200.fl import FederatedDataset, VQCFusedTrainer
    from qfedx_b200.models import VQC
    clients = [(torch.rand(256, 8) * math.pi, torch.randint(0, 3, (256,))) for _ in range(16)]
    data = FederatedDataset.from_clients(clients, dev)
    tr = VQCFusedTrainer(VQC(8, 6, 3).to(dev), data, epochs=1, batch_size=32, lr=0.05)
    for r in range(2):
        tr.train(tr.global_flat(), list(range(16)), r)
torch.cuda.synchronize()
print("done")
