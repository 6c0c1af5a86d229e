"""Sharded-statevector benchmark (BASELINE.json config 5), run under torchrun:

    torchrun --nproc-per-node 8 -m qfedx_b200.bench_sharded --qubits 30

Times (CUDA events, max over ranks) the fused global-qubit swap K6 - every rank pulls (W-1)/W of its
shard from its peers over NVLink and applies the following rotations in registers - against its NVLink
roofline, the NCCL all_to_all baseline B1 (+ a separate gate pass), and a full ansatz layer (gates/s).
"""
from __future__ import annotations

import argparse
import json
import math
import os

import numpy as np
import torch
import torch.distributed as dist

from .ops import _lib
from .ops.large import GATE, ansatz_gates, compile_passes, run_passes
from .parallel.sharded import ShardedState, nccl_swap_baseline, plan_sharded
from .utils.timing import max_over_ranks

NVLINK_MEASURED_GBS = 770.0   # peer copy per direction per GPU (B200_PROFILING.md)
NVLINK_NOMINAL_GBS = 900.0


def _time(fn, iters, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return max_over_ranks(s.elapsed_time(e) / iters)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=30)
    ap.add_argument("--layers", type=int, default=1)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n, g = args.qubits, int(math.log2(world))
    n_local = n - g
    st = ShardedState(n, world, rank, dev)
    st.reset_zero_state()
    theta = torch.randn(2 * n * args.layers, device=dev)
    ang = torch.rand(1, n, device=dev) * math.pi
    shard_bytes = (1 << n_local) * 8
    wire_bytes = shard_bytes * (world - 1) / world           # per GPU per direction

    # 1. fused swap + g rotations on the swapped-in qubits
    ops = torch.tensor([[GATE["rxrz"], b, -1, 2 * b, 2 * b + 1] for b in range(g)], dtype=torch.int32, device=dev)
    th = theta.reshape(1, -1)

    def fused():
        st.epoch += 1
        nxt = 1 - st.cur
        _lib.call("qfx_sv_global_swap_fused", _lib.ptr(st.tables[st.cur]), _lib.ptr(torch.view_as_real(st.bufs[nxt][0])),
                  _lib.c_int(n_local), _lib.c_int(g), _lib.c_int(rank), _lib.c_int(world), _lib.ptr(ops), _lib.c_int(g),
                  _lib.ptr(th), _lib.c_int(th.shape[1]), _lib.ptr(None), _lib.c_int(0), _lib.ptr(st.flag_table),
                  _lib.c_u32(st.epoch), _lib.stream_ptr())
        st.cur = nxt

    fused_ms = _time(fused, args.iters)

    # 2. baseline B1: NCCL all_to_all, then the same rotations as a separate local pass
    gate_pass = compile_passes(n_local, [(GATE["rx"], n_local - g + b, -1, 2 * b) for b in range(g)] +
                               [(GATE["rz"], n_local - g + b, -1, 2 * b + 1) for b in range(g)])

    def baseline():
        out = nccl_swap_baseline(st.shard(), world)
        run_passes(out[None], n_local, gate_pass, th)

    nccl_ms = _time(baseline, args.iters)

    def nccl_only():
        nccl_swap_baseline(st.shard(), world)

    nccl_only_ms = _time(nccl_only, args.iters)

    # 3. whole ansatz (feature map + L layers) through the planner: gates per second
    gates = ansatz_gates(n, args.layers)
    plan = plan_sharded(n, g, gates)

    def layer():
        st.run(plan, theta, ang)

    layer_ms = _time(layer, max(1, args.iters // 2), warmup=1)

    if rank == 0:
        roof_ms = wire_bytes / (NVLINK_MEASURED_GBS * 1e9) * 1e3
        print(json.dumps({
            "qubits": n, "world": world, "shard_gib": shard_bytes / 2 ** 30, "wire_gib_per_gpu": wire_bytes / 2 ** 30,
            "fused_swap_ms": fused_ms, "fused_swap_gbs": wire_bytes / fused_ms / 1e6,
            "roofline_ms_at_770": roof_ms, "fraction_of_measured_nvlink": roof_ms / fused_ms,
            "fraction_of_nominal_nvlink": wire_bytes / (NVLINK_NOMINAL_GBS * 1e9) * 1e3 / fused_ms,
            "nccl_all_to_all_plus_gate_pass_ms": nccl_ms, "nccl_all_to_all_only_ms": nccl_only_ms,
            "speedup_vs_nccl_path": nccl_ms / fused_ms,
            "ansatz": {"layers": args.layers, "gates": len(gates), "swaps": plan.n_swaps, "ms": layer_ms,
                       "gates_per_s": len(gates) / layer_ms * 1e3},
        }))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
