"""torchrun worker: K7 across GPUs (auto exchange: NVLS where available) vs the peer-store exchange vs NCCL, 1e8 floats.

usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_k7_multi.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from qfedx_b200.benchmarks import multi_gpu_aggregation  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    res = multi_gpu_aggregation(rank, world, dev)
    if rank == 0:
        print(json.dumps(res), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
