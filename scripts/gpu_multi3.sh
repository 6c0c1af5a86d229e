#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
Q=${2:-30}
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest multi rc=$?"; tail -5 gpurun_out/pytest_multi_n$N.log | cut -c1-600
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 -m qfedx_b200.bench_sharded --qubits $Q --layers 1 > gpurun_out/bench_sharded_n$N.log 2>&1; echo "bench_sharded rc=$?"; tail -2 gpurun_out/bench_sharded_n$N.log | cut -c1-1500
for cfg in vqc8 cfed; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config $cfg --steps 5 --warmup 3 --no-aux > gpurun_out/bench_${cfg}_n$N.log 2>&1; echo "bench $cfg N=$N rc=$?"; tail -1 gpurun_out/bench_${cfg}_n$N.log | cut -c1-400
done
