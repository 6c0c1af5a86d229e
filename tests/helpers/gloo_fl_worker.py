"""torchrun worker (CPU, gloo): a federated job split over ranks must equal the single-process job.

Each rank hosts a contiguous block of the clients (parallel/runtime.py: local_clients), trains them with the
torch trainer and aggregates through the all_reduce baseline aggregator; sampling and dropouts are decided over
the global population on every rank identically.  Rank 0 also runs the whole job alone and compares."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from qfedx_b200.fl import FederatedDataset, FederatedRunner, FLConfig, TorchLocalTrainer  # noqa: E402
from qfedx_b200.models import VQC  # noqa: E402
from qfedx_b200.parallel.runtime import client_owner, local_clients, make_aggregator  # noqa: E402


def clients(K=6, n=4):
    out = []
    for k in range(K):
        g = torch.Generator().manual_seed(100 + k)
        m = 10 + 3 * k                                   # unequal sizes -> unequal FedAvg weights
        y = torch.randint(0, 2, (m,), generator=g)
        x = torch.rand(m, n, generator=g) * 0.4 + y[:, None] * 1.8
        out.append((x, y))
    return out


def run(cl, gids, all_sizes, aggregator, rounds=3):
    torch.manual_seed(0)
    model = VQC(4, 2, 2)
    data = FederatedDataset.from_clients(cl, "cpu")
    # full-batch SGD: the result does not depend on the shuffle, so both layouts see the same arithmetic
    tr = TorchLocalTrainer(model, data, None, epochs=2, batch_size=64, lr=0.1, optimizer="sgd", momentum=0.0,
                           n_wrap=model.n_angle_params)
    cfg = FLConfig(num_rounds=rounds, mode="delta", clip=0.5, sigma=0.0, client_fraction=0.7, drop_prob=0.2, seed=3,
                   quiet=True)
    runner = FederatedRunner(tr, cfg, None, aggregator=aggregator, client_gids=gids, global_weights=all_sizes)
    for _ in range(rounds):
        runner.run_round()
    return runner.global_flat.clone()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    cl = clients()
    K = len(cl)
    all_sizes = torch.tensor([len(c[0]) for c in cl], dtype=torch.float32)
    mine = local_clients(rank, K, world)
    assert all(client_owner(k, K, world) == rank for k in mine)
    dim = sum(p.numel() for p in VQC(4, 2, 2).state_dict().values())
    agg = make_aggregator(rank, world, mine, all_sizes, dim, torch.device("cpu"), kind="nccl")
    split = run([cl[k] for k in mine], mine, all_sizes, agg)
    gathered = [torch.empty_like(split) for _ in range(world)]
    dist.all_gather(gathered, split)
    ok = all(torch.equal(gathered[0], t) for t in gathered)          # every rank holds the same global model
    if rank == 0:
        whole = run(cl, None, None, None)
        err = float((whole - split).abs().max())
        moved = float((whole - run(cl, None, None, None, rounds=0)).abs().max())
        ok = ok and err < 1e-5 and moved > 1e-3
        print(f"GLOO_FL_{'OK' if ok else 'FAIL'} err={err:.2e} moved={moved:.2e}", flush=True)
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
