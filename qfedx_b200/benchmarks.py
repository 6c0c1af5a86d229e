"""Secondary measurements reported under ``aux`` in bench.py's JSON line and runnable on their own
(``python -m qfedx_b200.benchmarks``): statevector gate throughput, kernel-vs-torch-baseline
ratios, aggregation latency.  Everything is timed with CUDA events after warm-up."""
from __future__ import annotations

import json
import math

import torch


def _time_ms(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def small_circuit_throughput(n=8, L=6, B=1 << 16):
    """Gates/s of the register-tiled forward (K1+K3) and forward+adjoint (K4a)."""
    from .ops import statevector as sv
    dev = torch.device("cuda")
    ang = torch.rand(B, n, device=dev) * math.pi
    theta = torch.randn(1, L, n, 2, device=dev)
    cot = torch.randn(B, n, device=dev)
    gates = B * (n + L * (2 * n + (n if n > 2 else n - 1)))
    fwd = _time_ms(lambda: sv.sv_forward(ang, theta))
    bwd = _time_ms(lambda: sv.sv_backward(ang, theta, cot))
    return {"n_qubits": n, "layers": L, "batch": B, "fwd_ms": fwd, "fwd_gates_per_s": gates / fwd * 1e3,
            "fwd_bwd_ms": bwd, "fwd_bwd_circuits_per_s": B / bwd * 1e3}


def torch_baseline_ratio(n=8, L=6, B=4096):
    """B1: the plain-torch complex statevector VQC (ops/ref.py) on the same GPU, fwd+bwd."""
    from .ops import ref
    from .ops import statevector as sv
    dev = torch.device("cuda")
    ang = torch.rand(B, n, device=dev) * math.pi
    theta = torch.randn(1, L, n, 2, device=dev, requires_grad=True)
    cot = torch.randn(B, n, device=dev)

    def base():
        theta.grad = None
        (ref.vqc_expvals_ref(ang, theta) * cot).sum().backward()

    ours = _time_ms(lambda: sv.sv_backward(ang, theta.detach(), cot), iters=20)
    torch_ms = _time_ms(base, iters=3, warmup=1)
    return {"batch": B, "ours_ms": ours, "torch_ms": torch_ms, "speedup": torch_ms / ours}


def aggregation_latency(K=64, D=102, secagg=False):
    from .ops.aggregate import fedavg_aggregate
    dev = torch.device("cuda")
    cp = torch.randn(K, D, device=dev)
    g = torch.randn(D, device=dev)
    w = torch.rand(K, device=dev) + 1
    out = torch.empty_like(g)
    scratch = torch.empty(K, device=dev)
    ms = _time_ms(lambda: fedavg_aggregate(cp, g, w, mode="delta", clip=1.0, sigma=0.5, secagg=secagg, out=out,
                                           scratch=scratch), iters=50)
    return {"K": K, "D": D, "secagg": secagg, "ms": ms, "bytes": K * D * 4, "gbps": K * D * 4 / ms / 1e6}


def aux_measurements(args=None, rank=0, world=1, local=0, skip=None):
    out = {}
    if rank != 0:
        return out
    out["sv8_throughput"] = small_circuit_throughput(8, 6, 1 << 15)
    out["sv4_throughput"] = small_circuit_throughput(4, 2, 1 << 18)
    out["vs_torch_baseline_8q"] = torch_baseline_ratio(8, 6, 2048)
    out["agg_vqc"] = aggregation_latency(64, 102)
    out["agg_tinycnn_dp_secagg"] = aggregation_latency(8, 113859, True)
    out["agg_1e8"] = aggregation_latency(4, 25_000_000)
    return out


if __name__ == "__main__":
    print(json.dumps(aux_measurements(), indent=1))
