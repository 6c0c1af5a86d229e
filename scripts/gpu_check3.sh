#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for mode in latency throughput; do for cl in 0 8; do
QFX_LANE_MODE=$mode QFX_CLUSTER=$cl timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8_${mode}_$cl.log 2>&1; echo "bench vqc8 $mode cluster=$cl rc=$?"; python -c "
import json,sys
l=open('gpurun_out/bench_vqc8_${mode}_$cl.log').read().strip().splitlines()[-1]
try:
    d=json.loads(l); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['final_eval'])
except Exception as e: print(l[-2000:])
"
done; done
QFX_LANE_MODE=latency timeout 600 python bench.py --config vqc4 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc4.log 2>&1; echo "bench vqc4 rc=$?"; tail -1 gpurun_out/bench_vqc4.log | cut -c1-200
QFX_LANE_MODE=throughput timeout 600 python bench.py --config vqc4 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc4_t.log 2>&1; echo "bench vqc4 thr rc=$?"; tail -1 gpurun_out/bench_vqc4_t.log | cut -c1-200
