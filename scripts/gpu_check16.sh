#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tinycnn.py tests/test_gpu_tinycnn_umma.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-300
for k in 1 8; do
QFX_BENCH_CLIENTS=$k timeout 600 python bench.py --config cfed --steps 4 --warmup 3 --no-aux > gpurun_out/bench_cfed_k$k.log 2>&1; echo "K=$k rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_cfed_k$k.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['final_eval'])" || tail -5 gpurun_out/bench_cfed_k$k.log
done
timeout 300 python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 45 --csv --log-file gpurun_out/launches_cfed.csv python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/ncu_cfed.log 2>&1; echo "ncu rc=$?"
