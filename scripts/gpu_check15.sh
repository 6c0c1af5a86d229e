#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_tinycnn_umma.py -m gpu -x -q > gpurun_out/pytest_umma.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_umma.log | cut -c1-300
