"""torchrun worker: NVLink probes (ring pull, all-to-all pull / push) and the fused swap on the same buffers.
usage: torchrun --nproc-per-node N scripts/p2p_probe.py [log2 of amplitudes per shard, default 26]"""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from qfedx_b200.ops import _lib
from qfedx_b200.parallel.symm import SymmetricBuffer
from qfedx_b200.utils.timing import max_over_ranks


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n_local = int(sys.argv[1]) if len(sys.argv) > 1 else 26
    nbytes = (1 << n_local) * 8
    src, dst = SymmetricBuffer(nbytes, rank, world, dev), SymmetricBuffer(nbytes, rank, world, dev)
    st, dt = src.peer_table(), dst.peer_table()
    rows = []
    for mode, name in ((0, "ring_pull"), (1, "a2a_pull"), (2, "a2a_push")):
        for unroll in (1, 2, 4, 8):
            for cps in (2, 4, 8):
                def run():
                    _lib.call("qfx_p2p_probe", _lib.ptr(st), _lib.ptr(dt), ctypes.c_size_t(nbytes), _lib.c_int(rank),
                              _lib.c_int(world), _lib.c_int(mode), _lib.c_int(unroll), _lib.c_int(cps), _lib.stream_ptr())
                for _ in range(2):
                    run()
                torch.cuda.synchronize(); dist.barrier()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(5):
                    run()
                e.record(); torch.cuda.synchronize()
                ms = max_over_ranks(s.elapsed_time(e) / 5)
                wire = nbytes if mode == 0 else nbytes * (world - 1) / world
                rows.append({"mode": name, "unroll": unroll, "ctas_per_sm": cps, "ms": round(ms, 4),
                             "gbs_per_dir": round(wire / ms / 1e6, 1)})
                dist.barrier()
    if rank == 0:
        best = {}
        for r in rows:
            if r["mode"] not in best or r["gbs_per_dir"] > best[r["mode"]]["gbs_per_dir"]:
                best[r["mode"]] = r
        print(json.dumps({"world": world, "shard_mib": nbytes >> 20, "best": best, "rows": rows}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
