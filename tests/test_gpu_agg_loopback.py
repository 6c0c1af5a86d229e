"""K7's multi-GPU phases on ONE GPU: every rank of a W-rank job runs as blockIdx.y of a single cooperative launch
(qfx_fedavg_fused_loopback), with all receive slots and flag words in this device's memory.  Covers what a 1-GPU box
otherwise never executes: publishing partial sums, the system-scope flag barrier, the slot reduction of the one-shot
exchange, and the reduce-scatter + all-gather of the two-shot exchange - against the host reference of the whole job."""
import ctypes

import numpy as np
import pytest
import torch

from qfedx_b200.fl.secagg import KeyAgreement
from qfedx_b200.ops import _lib
from qfedx_b200.ops.aggregate import MODES, _reference, pair_keys_tensor
from qfedx_b200.parallel.runtime import local_clients

pytestmark = pytest.mark.gpu


def loopback_aggregate(world, cp_all, glob, w, part, two_shot, epochs=2, pair_keys=None, **kw):
    dev = torch.device("cuda")
    K, D = cp_all.shape
    g = glob.to(dev).contiguous()
    wd = w.to(dev, torch.float32).contiguous()
    pd = part.to(dev, torch.int32).contiguous()
    rows = [[k for k in local_clients(r, K, world) if int(part[k])] for r in range(world)]
    cps = [cp_all[rw].to(dev).contiguous() if rw else None for rw in rows]
    gids = [torch.tensor(rw, dtype=torch.int32, device=dev) for rw in rows]
    outs = [torch.empty(D, device=dev) for _ in range(world)]
    scratch = [torch.zeros(max(1, len(rw)), device=dev) for rw in rows]
    chunk = ((D + world - 1) // world + 3) // 4 * 4
    words = (world + 1) * chunk if two_shot else world * D
    slot_sets = [[torch.zeros(words, dtype=torch.int64, device=dev) for _ in range(world)] for _ in range(2)]
    tables = [torch.tensor([s.data_ptr() for s in ss], dtype=torch.int64, device=dev) for ss in slot_sets]
    flags = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(world)]
    ftab = torch.tensor([f.data_ptr() for f in flags], dtype=torch.int64, device=dev)
    args_dev = torch.empty(world * 512, dtype=torch.uint8, device=dev)
    pk = pair_keys_tensor(pair_keys, dev) if pair_keys is not None else None
    P = ctypes.c_void_p * world
    for epoch in range(1, epochs + 1):          # several epochs: monotonic flags, alternating slot sets
        _lib.call("qfx_fedavg_fused_loopback", _lib.c_int(world),
                  P(*[c.data_ptr() if c is not None else None for c in cps]), P(*[t.data_ptr() for t in gids]),
                  (ctypes.c_int * world)(*[len(rw) for rw in rows]), P(*[o.data_ptr() for o in outs]),
                  P(*[s.data_ptr() for s in scratch]), _lib.ptr(g), _lib.ptr(wd), _lib.ptr(pd), _lib.c_int(K), _lib.c_int(D),
                  _lib.c_int(MODES[kw["mode"]]), _lib.c_int(kw["n_wrap"]), _lib.c_float(kw["clip"]), _lib.c_float(0.0),
                  _lib.c_float(kw["server_lr"]), _lib.c_u64(kw["seed"]), _lib.c_int(kw["rnd"]),
                  _lib.c_int(1 if kw["secagg"] else 0), _lib.ptr(pk), _lib.ptr(tables[epoch & 1]), _lib.ptr(ftab),
                  _lib.c_u32(epoch), _lib.c_int(1 if two_shot else 0), _lib.ptr(args_dev), _lib.stream_ptr())
    torch.cuda.synchronize()
    return outs


@pytest.mark.parametrize("two_shot", [False, True])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("K,D,secagg,clip,mode,dh", [(16, 4096, False, 0.7, "delta", False), (9, 1030, False, 0.0, "weights", False),
                                                   (8, 5000, True, 0.7, "delta", False), (8, 113859, False, 0.7, "delta", False),
                                                   (6, 2048, True, 0.5, "delta", True)])
def test_loopback_world_matches_host_reference(world, two_shot, K, D, secagg, clip, mode, dh):
    g = torch.Generator().manual_seed(1000 + K + D + world)
    glob = torch.randn(D, generator=g)
    cp_all = glob + 0.05 * torch.randn(K, D, generator=g) * (1 + torch.arange(K)[:, None])
    w = torch.randint(1, 500, (K,), generator=g).float()
    part = (torch.rand(K, generator=g) > 0.25).int()
    part[0] = 1
    keys = KeyAgreement(K, 5).table() if dh else None
    kw = dict(mode=mode, n_wrap=D // 2 if mode == "delta" else 0, clip=clip, server_lr=0.8, seed=5, rnd=3, secagg=secagg)
    outs = loopback_aggregate(world, cp_all, glob, w, part, two_shot, pair_keys=keys, **kw)
    ref = _reference(cp_all, glob, w, part, mode, kw["n_wrap"], clip, 0.0, 0.8, 5, 3, secagg, False, pair_keys=keys).new_global
    for r in range(world):
        assert torch.equal(outs[r], outs[0]), f"rank {r} disagrees with rank 0"       # bit-identical on every rank
    assert torch.allclose(outs[0].cpu(), ref, atol=3e-5), float((outs[0].cpu() - ref).abs().max())
