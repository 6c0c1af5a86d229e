#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/pytest_noalloc_alone.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu_full.log 2>&1; echo "full rc=$?"; tail -8 gpurun_out/pytest_gpu_full.log | cut -c1-300
