#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -30 gpurun_out/pytest_multi.log | cut -c1-400
for cfg in vqc8 cfed; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config $cfg --steps 5 --warmup 3 --no-aux > gpurun_out/bench_${cfg}_n$N.log 2>&1; echo "bench $cfg N=$N rc=$?"; tail -2 gpurun_out/bench_${cfg}_n$N.log | cut -c1-600
done
