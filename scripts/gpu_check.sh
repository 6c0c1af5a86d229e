#!/bin/bash
# First-contact GPU script: tests, smoke, short benches.  Each step under its own timeout.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8.log 2>&1; echo "bench vqc8 rc=$?"; tail -3 gpurun_out/bench_vqc8.log
timeout 600 python bench.py --config vqc4 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc4.log 2>&1; echo "bench vqc4 rc=$?"; tail -3 gpurun_out/bench_vqc4.log
timeout 600 python -m qfedx_b200.benchmarks > gpurun_out/aux.log 2>&1; echo "aux rc=$?"; tail -60 gpurun_out/aux.log
