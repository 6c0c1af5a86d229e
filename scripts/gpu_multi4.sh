#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
for cfg in cfed vqc8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --config $cfg --steps 5 --warmup 3 --no-aux > gpurun_out/bench_${cfg}_n$N.log 2>&1; echo "bench $cfg N=$N rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_${cfg}_n$N.log').read().strip().splitlines()[-1]); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])" || tail -5 gpurun_out/bench_${cfg}_n$N.log
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.log 2>&1; echo "ref N=$N rc=$?"; tail -1 gpurun_out/bench_ref_n$N.log | cut -c1-300
