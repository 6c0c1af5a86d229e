#!/bin/bash
# One-GPU validation on the B200 box: gpurun --timeout 3000 -- 'bash scripts/gpu_tests.sh'
# full GPU test suite, smoke(), the default benchmark (both arms) - logs under gpurun_out/.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu_full.log 2>&1; echo "full rc=$?"; tail -4 gpurun_out/pytest_gpu_full.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log | cut -c1-300
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.log 2>&1; echo "bench default rc=$?"; tail -1 gpurun_out/bench_default.log | cut -c1-3000
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "bench ref rc=$?"; tail -1 gpurun_out/bench_ref.log | cut -c1-600
timeout 300 python bench.py --impl reference --config vqc8 > gpurun_out/bench_ref_vqc8.log 2>&1; echo "bench ref vqc8 rc=$?"; tail -1 gpurun_out/bench_ref_vqc8.log | cut -c1-400
