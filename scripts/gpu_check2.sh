#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --config vqc8 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc8.log 2>&1; echo "bench vqc8 rc=$?"; tail -1 gpurun_out/bench_vqc8.log
timeout 600 python bench.py --config vqc4 --steps 5 --warmup 3 --no-aux > gpurun_out/bench_vqc4.log 2>&1; echo "bench vqc4 rc=$?"; tail -1 gpurun_out/bench_vqc4.log
timeout 600 python -m qfedx_b200.benchmarks > gpurun_out/aux.log 2>&1; echo "aux rc=$?"; tail -70 gpurun_out/aux.log
timeout 300 python bench.py --config vqc8 --steps 2 --warmup 1 --no-aux > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vqc_local_train -c 1 -o gpurun_out/prof_vqc_train python bench.py --config vqc8 --steps 2 --warmup 1 --no-aux > gpurun_out/ncu_train.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_train.log
