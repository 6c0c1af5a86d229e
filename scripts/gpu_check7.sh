#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tinycnn.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --config cfed --steps 5 --warmup 3 --no-aux > gpurun_out/bench_cfed.log 2>&1; echo "bench cfed rc=$?"; tail -1 gpurun_out/bench_cfed.log | cut -c1-300
timeout 300 python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 45 --csv --log-file gpurun_out/launches_cfed.csv python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/ncu_cfed.log 2>&1; echo "ncu rc=$?"
