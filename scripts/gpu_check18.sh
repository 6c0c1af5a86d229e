#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train_agg.py tests/test_gpu_statevector.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-300
for cfg in vqc8 vqc4; do
timeout 600 python bench.py --config $cfg --steps 5 --warmup 3 --no-aux > gpurun_out/bench_$cfg.log 2>&1; echo "$cfg rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_$cfg.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['final_eval'], d['clocks'])"
done
QFX_LANE_MODE=throughput timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8_t.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench_vqc8_t.log').read().strip().splitlines()[-1]); print('throughput mode', d['ms_per_step'], d['value'])"
