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


def nvlink_probes(rank, world, dev, n_local, iters=5):
    """Peer bandwidth measured in this job on these GPUs (csrc/p2p_probe.cu): ring pull (every rank copies its
    neighbour's shard: all links busy in both directions), all-to-all pull (the swap's pattern without gates) and
    all-to-all push.  GB/s per direction per GPU, max time over ranks."""
    import ctypes
    from .parallel.symm import SymmetricBuffer
    nbytes = (1 << n_local) * 8
    src, dst = SymmetricBuffer(nbytes, rank, world, dev), SymmetricBuffer(nbytes, rank, world, dev)
    stab, dtab = src.peer_table(), dst.peer_table()
    out = {}
    for mode, name in ((0, "ring_pull"), (1, "all_to_all_pull"), (2, "all_to_all_push")):
        best = None
        for unroll, cps in ((4, 8), (2, 4)):
            def run():
                _lib.call("qfx_p2p_probe", _lib.ptr(stab), _lib.ptr(dtab), ctypes.c_size_t(nbytes), _lib.c_int(rank),
                          _lib.c_int(world), _lib.c_int(mode), _lib.c_int(unroll), _lib.c_int(cps), _lib.stream_ptr())
            ms = _time(run, iters)
            best = ms if best is None else min(best, ms)
        wire = nbytes if mode == 0 else nbytes * (world - 1) / world
        out[name + "_gbs"] = wire / best / 1e6
    del src, dst
    return out


def swap_measurements(rank, world, dev, n, layers=1, iters=5, with_layer=True, with_probes=True):
    """K6 against its rooflines and baselines; every rank calls this, the dict is complete on every rank."""
    g = int(math.log2(world))
    n_local = n - g
    st = ShardedState(n, world, rank, dev)
    st.reset_zero_state()
    theta = torch.randn(2 * n * layers, device=dev)
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

    by_cps = {}
    for cps in (4, 8, 12):
        _lib.call("qfx_sv_swap_set_ctas_per_sm", _lib.c_int(cps), launches=0)
        by_cps[cps] = _time(fused, iters)
    best_cps = min(by_cps, key=by_cps.get)
    _lib.call("qfx_sv_swap_set_ctas_per_sm", _lib.c_int(best_cps), launches=0)
    fused_ms = by_cps[best_cps]

    # 2. baseline B1: NCCL all_to_all, then the same rotations as a separate local pass
    gate_pass = compile_passes(n_local, [(GATE["rx"], n_local - g + b, -1, 2 * b) for b in range(g)] +
                               [(GATE["rz"], n_local - g + b, -1, 2 * b + 1) for b in range(g)])

    def baseline():
        out = nccl_swap_baseline(st.shard(), world)
        run_passes(out[None], n_local, gate_pass, th)

    nccl_ms = _time(baseline, iters)

    def nccl_only():
        nccl_swap_baseline(st.shard(), world)

    nccl_only_ms = _time(nccl_only, iters)
    res = {
        "qubits": n, "world": world, "shard_gib": shard_bytes / 2 ** 30, "wire_gib_per_gpu": wire_bytes / 2 ** 30,
        "fused_swap_ms": fused_ms, "fused_swap_gbs": wire_bytes / fused_ms / 1e6, "ctas_per_sm": best_cps,
        "fused_swap_ms_by_ctas_per_sm": by_cps,
        "fraction_of_nominal_nvlink_900": wire_bytes / (NVLINK_NOMINAL_GBS * 1e9) * 1e3 / fused_ms,
        "fraction_of_quoted_peer_copy_770": wire_bytes / (NVLINK_MEASURED_GBS * 1e9) * 1e3 / fused_ms,
        "nccl_all_to_all_plus_gate_pass_ms": nccl_ms, "nccl_all_to_all_only_ms": nccl_only_ms,
        "nccl_all_to_all_gbs": wire_bytes / nccl_only_ms / 1e6,
        "speedup_vs_nccl_path": nccl_ms / fused_ms,
    }
    if with_layer:
        # 3. whole ansatz (feature map + L layers) through the planner: gates per second
        gates = ansatz_gates(n, layers)
        plan = plan_sharded(n, g, gates)
        layer_ms = _time(lambda: st.run(plan, theta, ang), max(1, iters // 2), warmup=1)
        res["ansatz"] = {"layers": layers, "gates": len(gates), "swaps": plan.n_swaps, "ms": layer_ms,
                         "gates_per_s": len(gates) / layer_ms * 1e3}
    del st
    torch.cuda.empty_cache()
    if with_probes:
        res["measured_in_this_run"] = nvlink_probes(rank, world, dev, n_local, iters)
        ring = res["measured_in_this_run"]["ring_pull_gbs"]
        res["fraction_of_measured_ring_pull"] = res["fused_swap_gbs"] / ring
    return res


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
    res = swap_measurements(rank, world, dev, args.qubits, args.layers, args.iters)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
