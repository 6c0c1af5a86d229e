"""torchrun worker: fused cross-GPU aggregation (K7, peer stores over NVLink) vs the NCCL baseline."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from qfedx_b200.parallel.runtime import MultiGpuAggregator, NcclAggregator, local_clients  # noqa: E402


def check(rank, world, dev, K, D, secagg, clip, exchange):
    ok = True
    g = torch.Generator().manual_seed(1234 + K + D)                  # same stream on every rank
    glob = torch.randn(D, generator=g)
    cp_all = glob + 0.05 * torch.randn(K, D, generator=g) * (1 + torch.arange(K)[:, None])
    w = torch.randint(1, 500, (K,), generator=g).float()
    part = (torch.rand(K, generator=g) > 0.25).int()
    part[0] = 1
    mine = local_clients(rank, K, world)
    rows = [k for k in mine if part[k]]
    cp = cp_all[rows].to(dev) if rows else None
    fused = MultiGpuAggregator(rank, world, D, dev, exchange=exchange)
    base = NcclAggregator(rank, world, D, dev)
    b = None
    for rnd in range(3):                                                 # several epochs: flags, slot parity
        a = fused(cp, glob.to(dev), w, part, mode="delta", n_wrap=D // 2, clip=clip, server_lr=0.8, seed=5,
                  rnd=rnd, secagg=secagg, gids=rows).new_global
        b = base(cp, glob.to(dev), w, part, mode="delta", n_wrap=D // 2, clip=clip, server_lr=0.8, seed=5,
                 rnd=rnd, secagg=False, gids=rows).new_global
        torch.cuda.synchronize()
        err = float((a - b).abs().max())
        gathered = [torch.empty_like(a) for _ in range(world)]            # every rank must hold the same result
        dist.all_gather(gathered, a)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        if err > 3e-5 or not same:
            ok = False
            print(f"rank {rank}: K={K} D={D} secagg={secagg} exchange={exchange} round {rnd}: err={err:.3e} identical={same}", flush=True)
    # back-to-back rounds with every rank delayed by a different random amount before each launch: a fast
    # rank runs ahead into the next epoch while a slow one is still reducing the previous slots
    skew = torch.Generator().manual_seed(99 + rank)
    outs = []
    for rnd in range(3, 11):
        torch.cuda._sleep(int(torch.randint(0, 3_000_000, (1,), generator=skew)))
        outs.append(fused(cp, glob.to(dev), w, part, mode="delta", n_wrap=D // 2, clip=clip, server_lr=0.8,
                          seed=5, rnd=rnd, secagg=secagg, gids=rows).new_global.clone())
    torch.cuda.synchronize()
    for i, a in enumerate(outs):
        err = float((a - b).abs().max())
        if err > 3e-5:
            ok = False
            print(f"rank {rank}: K={K} D={D} secagg={secagg} exchange={exchange} skewed round {i}: err={err:.3e}", flush=True)
    del fused
    return ok


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    ok = True
    from qfedx_b200.parallel.symm import MulticastBuffer
    sup = torch.tensor([1 if MulticastBuffer.supported(dev) else 0], device=dev, dtype=torch.int32)
    dist.all_reduce(sup, op=dist.ReduceOp.MIN)
    exchanges = ("one_shot", "two_shot") + (("nvls",) if int(sup) else ())
    if rank == 0:
        print("exchanges under test:", exchanges, flush=True)
    for K, D, secagg, clip in [(16, 102, False, 0.0), (16, 4096, False, 0.7), (16, 113859, False, 0.7), (8, 5000, True, 0.7), (5, 300, True, 0.0)]:
        for exchange in exchanges:                                       # every exchange algorithm the box supports
            ok = check(rank, world, dev, K, D, secagg, clip, exchange) and ok
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MULTIGPU_AGG_OK" if int(flag) == 1 else "MULTIGPU_AGG_FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
