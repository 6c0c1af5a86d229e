#!/bin/bash
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_unitary_gemm.py -m gpu -x -q > gpurun_out/pytest_k2.log 2>&1; echo "pytest k2 rc=$?" | tee -a gpurun_out/pytest_k2.log
tail -40 gpurun_out/pytest_k2.log
nvidia-smi --query-gpu=name,memory.used --format=csv
timeout 300 python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv2_ -s 30 -c 3 -o gpurun_out/prof_conv2 python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/ncu_conv2.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_conv2.log | cut -c1-300
