#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_large.py -m gpu -x -q --timeout 120 > gpurun_out/pytest_large.log 2>&1; echo "pytest large rc=$?"
tail -40 gpurun_out/pytest_large.log | cut -c1-300
