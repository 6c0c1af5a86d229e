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


def aggregation_latency(K=64, D=102, secagg=False, mode="delta", clip=1.0, sigma=0.5, iters=50):
    """K7 latency and the HBM traffic it implies: weights mode reads K*D once; delta mode with a clip reads the
    client rows twice (norm sweep, then the weighted sum) plus the global vector twice, and writes D."""
    from .ops.aggregate import fedavg_aggregate
    dev = torch.device("cuda")
    cp = torch.randn(K, D, device=dev)
    g = torch.randn(D, device=dev)
    w = torch.rand(K, device=dev) + 1
    out = torch.empty_like(g)
    scratch = torch.empty(K, device=dev)
    ms = _time_ms(lambda: fedavg_aggregate(cp, g, w, mode=mode, clip=clip, sigma=sigma, secagg=secagg, out=out,
                                           scratch=scratch), iters=iters)
    sweeps = 2 if (mode == "delta" and clip > 0) else 1
    traffic = 4 * (sweeps * K * D + (sweeps if mode == "delta" else 0) * D + D)
    return {"K": K, "D": D, "secagg": secagg, "mode": mode, "clip": clip, "sigma": sigma, "ms": ms,
            "bytes": K * D * 4, "hbm_traffic_bytes": traffic, "hbm_gbps": traffic / ms / 1e6}


def unitary_gemm_throughput(n=8, L=6, B=1 << 15):
    """K2: samples/s and TFLOP/s of the tcgen05 complex GEMM (3xTF32 and plain TF32) vs the register kernel."""
    from .ops import statevector as sv
    from .ops.unitary_gemm import UnitaryGemm
    dev = torch.device("cuda")
    theta = torch.randn(1, L, n, 2, device=dev)
    ang = torch.rand(B, n, device=dev) * math.pi
    out = {"n_qubits": n, "batch": B}
    m = 2 << n
    for prec in (3, 1):
        ug = UnitaryGemm(theta, precision=prec)
        bp = ug._pad(B)
        b_hi = torch.empty(bp, m, device=dev); b_lo = torch.empty_like(b_hi)
        from .ops import _lib
        from .ops.statevector import BASIS
        _lib.call("qfx_unitary_pack_product_states", _lib.ptr(ang), _lib.c_int(n), _lib.c_int(BASIS["ry"]), _lib.c_int(B),
                  _lib.c_int(bp), _lib.ptr(b_hi), _lib.ptr(b_lo), _lib.stream_ptr())
        gemm_only = _time_ms(lambda: ug._gemm(b_hi, b_lo, bp, True, False), iters=10)
        e2e = _time_ms(lambda: ug.expvals(ang), iters=10)
        flops = 2.0 * m * m * bp * prec
        out[f"tf32x{prec}"] = {"gemm_ms": gemm_only, "gemm_tflops": flops / gemm_only / 1e9,
                               "expvals_pipeline_ms": e2e, "samples_per_s": B / e2e * 1e3}
    reg = _time_ms(lambda: sv.sv_forward(ang, theta), iters=10)
    out["register_kernel_ms"] = reg
    return out


def unitary_gemm_vs_engines():
    """Where K2 earns its place in the product (models/vqc.py: inference_expvals picks it there): deep circuits over
    large batches - at 12 qubits (no register kernel exists) against the pass engine K5, at 10 qubits against the register
    kernel K1.  Both competitors grow with the depth, the GEMM does not; the shallow 12-qubit row shows the other side of
    the break-even (the selection rule does not pick the GEMM there).  Times include building and
    packing the unitary (once per parameter set) separately from the per-batch pipeline."""
    from .models import VQC
    dev = torch.device("cuda")
    out = {}
    shapes = {"12q_16L_4096": (12, 16, 4096, "passes"), "12q_4L_4096": (12, 4, 4096, "passes"),
              "10q_24L_16384": (10, 24, 16384, "registers")}
    from .ops.unitary_gemm import prefers_unitary_gemm
    for name, (n, L, B, other) in shapes.items():
        torch.manual_seed(0)
        m = VQC(n, L, 2).to(dev)
        x = torch.rand(B, n, device=dev) * math.pi
        t_build = _time_ms(lambda: (setattr(m, "_ug_key", None), m.inference_expvals(x[:64], "gemm")), iters=3, warmup=1)
        a = m.inference_expvals(x, "gemm")
        b = m.inference_expvals(x, other)
        t_gemm = _time_ms(lambda: m.inference_expvals(x, "gemm"), iters=10)
        t_other = _time_ms(lambda: m.inference_expvals(x, other), iters=5)
        out[name] = {"n_qubits": n, "layers": L, "batch": B, "gemm_ms": t_gemm, "unitary_build_and_pack_ms": t_build,
                     "other_engine": other, "other_ms": t_other, "speedup": t_other / t_gemm,
                     "max_abs_diff": float((a - b).abs().max()), "auto_picks_gemm": prefers_unitary_gemm(n, L, B)}
    return out


def large_state_throughput(n=20, L=2, batch=16, tile_bits=12, low_bits=3, register_engine=True):
    """K5: gates/s of the pass engine on `batch` n-qubit states (config 4 shape: 16 clients x 20 qubits)."""
    from .ops.large import LargeVQC, expvals_z
    dev = torch.device("cuda")
    vqc = LargeVQC(n, L, tile_bits=tile_bits, low_bits=low_bits, register_engine=register_engine)
    ang = torch.rand(batch, n, device=dev) * math.pi
    theta = torch.randn(1, vqc.n_params, device=dev)
    ms = _time_ms(lambda: vqc.states(ang, theta), iters=5, warmup=2)
    st = vqc.states(ang, theta)
    ez = _time_ms(lambda: expvals_z(st, n), iters=5, warmup=1)
    state_bytes = (1 << n) * 8 * batch
    return {"n_qubits": n, "layers": L, "batch": batch, "tile_bits": tile_bits, "low_bits": low_bits, "register_engine": register_engine,
            "passes": len(vqc.passes), "gates": vqc.n_gates, "ms": ms,
            "gates_per_s": vqc.n_gates * batch / ms * 1e3,
            "effective_gbs": 2 * state_bytes * len(vqc.passes) / ms / 1e6, "expz_ms": ez,
            # what training uses: all <Z_q> of the circuit, final ring folded into the parity readout
            "readout_passes": len(vqc.readout_passes),
            "expvals_ms": _time_ms(lambda: vqc.expvals(ang, theta), iters=5, warmup=2)}


def multi_gpu_aggregation(rank, world, dev, D=100_000_000 // 8 * 8, iters=5):
    """K7 across GPUs inside one cooperative launch (in-switch reduction where the job has a multicast object, else
    peer stores + loads) against NCCL all_reduce on a 1e8-float vector: every rank holds ONE client row, plain FedAvg
    of weights (one sweep), max time over ranks.  The peer-store exchange is timed as well when NVLS was picked."""
    import torch.distributed as dist
    from .parallel.runtime import MultiGpuAggregator
    from .utils.timing import max_over_ranks
    fused = MultiGpuAggregator(rank, world, D, dev)
    peer = MultiGpuAggregator(rank, world, D, dev, exchange="two_shot") if fused.exchange == "nvls" else None
    row = torch.randn(1, D, device=dev)
    glob = torch.zeros(D, device=dev)
    out = torch.empty(D, device=dev)
    fout = fused.result_row if fused.exchange == "nvls" else out      # NVLS: the broadcast row is the result (no copy)
    w = torch.ones(world)

    def run_fused():
        fused(row, glob, w, None, mode="weights", gids=[rank], out=fout)

    def run_peer():
        peer(row, glob, w, None, mode="weights", gids=[rank], out=out)

    def run_nccl():
        out.copy_(row[0])
        out.mul_(1.0 / world)
        dist.all_reduce(out)

    def run_nccl_only():
        dist.all_reduce(out)

    res = {"exchange": fused.exchange}
    cases = [("fused_ms", run_fused), ("nccl_all_reduce_ms", run_nccl), ("nccl_all_reduce_alone_ms", run_nccl_only)] + ([("fused_peer_two_shot_ms", run_peer)] if peer else [])
    for name, fn in cases:
        for _ in range(2):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record(); torch.cuda.synchronize()
        res[name] = max_over_ranks(s.elapsed_time(e) / iters)
    run_fused(); a = fout.clone(); run_nccl()
    res["max_abs_diff_vs_nccl"] = float((a - out).abs().max())
    res.update({"floats": D, "world": world,
                "fused_algbw_gbs": D * 4 / res["fused_ms"] / 1e6, "nccl_algbw_gbs": D * 4 / res["nccl_all_reduce_ms"] / 1e6,
                "fused_vs_nccl": res["nccl_all_reduce_ms"] / res["fused_ms"]})
    del fused, peer
    torch.cuda.empty_cache()
    return res


def multi_gpu_aux(rank, world, local):
    """Secondary measurements that need every rank: the K6 swap against its rooflines and baselines, and K7 across
    GPUs against NCCL.  Sized to the job: 27 local qubits per GPU (1 GiB shards) as in BASELINE config 5."""
    import math
    from .bench_sharded import swap_measurements
    dev = torch.device("cuda", local)
    out = {}
    g = int(math.log2(world))
    if (1 << g) == world and world in (2, 4, 8):
        out["k6_swap"] = swap_measurements(rank, world, dev, 27 + g, layers=1, iters=5, with_layer=(world == 8))
    out["k7_allreduce_1e8"] = multi_gpu_aggregation(rank, world, dev)
    return out


def aux_measurements(args=None, rank=0, world=1, local=0, skip=None):
    out = {}
    if world > 1:
        try:
            out.update(multi_gpu_aux(rank, world, local))
        except Exception as exc:   # keep the single-GPU numbers if a collective measurement fails
            out["multi_gpu_error"] = f"{type(exc).__name__}: {exc}"
    if rank != 0:
        return out

    def safe(key, fn, *a, **kw):     # one failing secondary number must not take the others with it
        try:
            out[key] = fn(*a, **kw)
        except Exception as exc:
            out[key] = {"error": f"{type(exc).__name__}: {exc}"}

    safe("sv8_throughput", small_circuit_throughput, 8, 6, 1 << 15)
    safe("sv4_throughput", small_circuit_throughput, 4, 2, 1 << 18)
    safe("vs_torch_baseline_8q", torch_baseline_ratio, 8, 6, 2048)
    safe("agg_vqc", aggregation_latency, 64, 102)
    safe("agg_tinycnn_dp_secagg", aggregation_latency, 8, 113859, True)
    safe("agg_1e8_dp", aggregation_latency, 4, 25_000_000, iters=10)
    safe("agg_1e8_clip", aggregation_latency, 4, 25_000_000, sigma=0.0, iters=10)
    safe("agg_1e8_weights", aggregation_latency, 4, 25_000_000, mode="weights", clip=0.0, sigma=0.0, iters=10)
    safe("unitary_gemm_8q", unitary_gemm_throughput, 8, 6, 1 << 15)
    safe("unitary_gemm_vs_engines", unitary_gemm_vs_engines)
    safe("sv20_pass_engine", large_state_throughput, 20, 2, 16)
    return out


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "tiles":
        res = []
        for tb, lb in [(12, 3), (12, 4), (12, 5), (11, 3), (10, 3), (12, 2)]:
            res.append(large_state_throughput(20, 2, 16, tb, lb))
            res.append(large_state_throughput(26, 1, 1, tb, lb))
        print(json.dumps(res, indent=1))
    else:
        print(json.dumps(aux_measurements(), indent=1))
