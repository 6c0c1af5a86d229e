#!/bin/bash
mkdir -p gpurun_out
for cl in 0 8 2; do
QFX_BENCH_CLIENTS=8 QFX_CLUSTER=$cl timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8_k8_c$cl.log 2>&1; echo "vqc8 8 clients cluster=$cl rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_vqc8_k8_c$cl.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['clocks'])"
done
