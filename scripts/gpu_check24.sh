#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train_agg.py tests/test_gpu_tinycnn.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
for cfg in vqc8 vqc4 cfed; do
timeout 600 python bench.py --config $cfg --steps 5 --warmup 3 --no-aux > gpurun_out/bench_$cfg.log 2>&1; echo "$cfg rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_$cfg.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['final_eval'])"
done
