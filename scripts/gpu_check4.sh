#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
QFX_LANE_MODE=latency timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8.log 2>&1; echo "bench vqc8 rc=$?"; tail -1 gpurun_out/bench_vqc8.log | cut -c1-250
timeout 900 python bench.py --config cfed --steps 3 --warmup 3 --no-aux > gpurun_out/bench_cfed.log 2>&1; echo "bench cfed rc=$?"; tail -3 gpurun_out/bench_cfed.log | cut -c1-1500
timeout 1200 python bench.py --config cfed --impl reference --steps 2 --warmup 1 > gpurun_out/bench_cfed_ref.log 2>&1; echo "bench cfed ref rc=$?"; tail -3 gpurun_out/bench_cfed_ref.log | cut -c1-1500
