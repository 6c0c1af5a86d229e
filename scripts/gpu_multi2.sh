#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
Q=${2:-27}
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -12 gpurun_out/pytest_multi.log | cut -c1-600
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 -m qfedx_b200.bench_sharded --qubits $Q --layers 1 > gpurun_out/bench_sharded_n$N.log 2>&1; echo "bench_sharded rc=$?"; tail -3 gpurun_out/bench_sharded_n$N.log | cut -c1-1500
