"""Runs each headline kernel a few times at its benchmark shape (for ncu captures)."""
import math
import sys

import torch

sys.path.insert(0, ".")
from qfedx_b200.ops import statevector as sv
from qfedx_b200.ops.aggregate import fedavg_aggregate
from qfedx_b200.ops.large import LargeVQC
from qfedx_b200.ops.unitary_gemm import UnitaryGemm
from qfedx_b200.parallel.sharded import ShardedState, plan_sharded
from qfedx_b200.ops.large import ansatz_gates

dev = torch.device("cuda")
torch.manual_seed(0)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "sv"):
    ang = torch.rand(1 << 16, 8, device=dev) * math.pi
    theta = torch.randn(1, 6, 8, 2, device=dev)
    cot = torch.randn(1 << 16, 8, device=dev)
    for _ in range(3):
        sv.sv_forward(ang, theta)
        sv.sv_backward(ang, theta, cot)
if which in ("all", "gemm"):
    ug = UnitaryGemm(torch.randn(1, 6, 8, 2, device=dev))
    ang = torch.rand(1 << 15, 8, device=dev) * math.pi
    for _ in range(3):
        ug.expvals(ang, want_states=True)
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
        fedavg_aggregate(cp, g, w, mode="delta", clip=1.0, sigma=0.5)
    for _ in range(3):
        fedavg_aggregate(cp, g, w, mode="weights")
if which in ("all", "large"):
    vqc = LargeVQC(20, 2)
    ang = torch.rand(16, 20, device=dev) * math.pi
    theta = torch.randn(1, vqc.n_params, device=dev)
    for _ in range(2):
        vqc.expvals(ang, theta)
if which in ("all", "swap"):
    n, world = 27, 2          # 2 x 512 MiB shards in one device's memory (loopback)
    plan = plan_sharded(n, 1, ansatz_gates(n, 1))
    st = ShardedState(n, world, loopback=True)
    st.reset_zero_state()
    st.run(plan, torch.randn(2 * n, device=dev), torch.rand(1, n, device=dev))
torch.cuda.synchronize()
print("done")
