"""torchrun worker: sharded statevector (K6, in-kernel peer loads) vs the single-GPU pass engine (K5)."""
import math
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from qfedx_b200.ops.large import LargeVQC, ansatz_gates  # noqa: E402
from qfedx_b200.parallel.sharded import ShardedState, plan_sharded  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    ok = True
    for n, L in [(18, 2), (20, 1)]:
        g = torch.Generator().manual_seed(n)
        ang = (torch.rand(1, n, generator=g) * math.pi).to(dev)
        theta = torch.randn(2 * n * L, generator=g).to(dev)
        plan = plan_sharded(n, int(math.log2(world)), ansatz_gates(n, L))
        st = ShardedState(n, world, rank, dev)
        for rep in range(2):                      # twice: buffer ping-pong and flag epochs keep working
            st.reset_zero_state()
            dist.barrier()
            st.run(plan, theta, ang)
            E = st.expvals(plan)
            E_ref = LargeVQC(n, L).expvals(ang, theta.reshape(1, -1))[0]
            err = float((E - E_ref).abs().max())
            if err > 5e-5:
                ok = False
                print(f"rank {rank}: n={n} L={L} rep {rep}: err {err:.3e}", flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MULTIGPU_SHARDED_OK" if int(flag) == 1 else "MULTIGPU_SHARDED_FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
