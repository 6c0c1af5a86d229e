#!/bin/bash
# Multi-GPU validation: gpurun --gpus 8 --timeout 1500 -- 'bash scripts/gpu_multi.sh 8'
# torchrun tests (fused aggregation, sharded statevector), the 30-qubit swap benchmark, bench.py at N GPUs.
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest multi rc=$?"; tail -3 gpurun_out/pytest_multi_n$N.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 -m qfedx_b200.bench_sharded --qubits 30 --layers 1 > gpurun_out/bench_sharded_n$N.log 2>&1; echo "bench_sharded rc=$?"; tail -1 gpurun_out/bench_sharded_n$N.log | cut -c1-1200
for cfg in cfed vqc8 vqc20 vqc30; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config $cfg --steps ${STEPS:-5} --warmup 3 --no-aux > gpurun_out/bench_${cfg}_n$N.log 2>&1; echo "bench $cfg N=$N rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_${cfg}_n$N.log').read().strip().splitlines()[-1]); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])" || tail -5 gpurun_out/bench_${cfg}_n$N.log
done
