"""ONE process, two GPUs: the fused swap kernel K6 reading a peer GPU's shard over NVLink - the workload of the ncu capture
that records NVLink byte counters for K6 (no torchrun, no kernel waits on another: rank 0's kernel runs on cuda:0, then
rank 1's on cuda:1, with the flag barrier switched off because the host orders the launches).

usage: python scripts/profile_swap_2gpu.py [local qubits per shard, default 25]"""
import ctypes
import math
import sys

import torch

sys.path.insert(0, ".")
from qfedx_b200.ops import _lib
from qfedx_b200.ops.large import GATE


def main():
    assert torch.cuda.device_count() >= 2, "needs two GPUs"
    n_local = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    world, g = 2, 1
    _lib.call("qfx_enable_peer", _lib.c_int(0), _lib.c_int(1), launches=0)
    _lib.call("qfx_enable_peer", _lib.c_int(1), _lib.c_int(0), launches=0)
    size = 1 << n_local
    src = [torch.randn(size, 2, device=f"cuda:{r}") for r in range(world)]
    dst = [torch.empty(size, 2, device=f"cuda:{r}") for r in range(world)]
    ref = [s.clone() for s in src]
    for r in range(world):
        with torch.cuda.device(r):
            table = torch.tensor([s.data_ptr() for s in src], dtype=torch.int64, device=f"cuda:{r}")
            ops = torch.tensor([[GATE["rxrz"], 0, -1, 0, 1]], dtype=torch.int32, device=f"cuda:{r}")
            th = torch.tensor([[0.3, -0.7]], device=f"cuda:{r}")
            for _ in range(3):
                _lib.call("qfx_sv_global_swap_fused", _lib.ptr(table), _lib.ptr(dst[r]), _lib.c_int(n_local), _lib.c_int(g),
                          _lib.c_int(r), _lib.c_int(world), _lib.ptr(ops), _lib.c_int(1), _lib.ptr(th), _lib.c_int(2),
                          _lib.ptr(None), _lib.c_int(0), _lib.ptr(None), _lib.c_u32(0),
                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
    # sanity: the exchange really read the peer (rank 0's upper half comes from rank 1's lower half before the gate mixes them)
    half = size // 2
    c, s_ = math.cos(0.15), math.sin(0.15)
    a0 = torch.view_as_complex(ref[0][:half].cpu().contiguous())          # slice 0 of rank 0
    a1 = torch.view_as_complex(ref[1][:half].cpu().contiguous())          # slice 0 of rank 1 (pulled over NVLink)
    cp, sp = math.cos(-0.35), math.sin(-0.35)
    m00 = complex(cp * c, -sp * c); m01 = complex(-sp * s_, -cp * s_)
    out0 = torch.view_as_complex(dst[0][:half].cpu().contiguous())
    err = float((out0 - (m00 * a0 + m01 * a1)).abs().max())
    print(f"wire bytes per rank and swap: {size * 8 // 2}; max |err| vs host = {err:.2e}")
    assert err < 1e-4
    print("done")


if __name__ == "__main__":
    main()
