#!/bin/bash
# Configs 4 and 5 through bench.py on one GPU (vqc30 at N=1 keeps the whole 8 GiB state on the GPU).
mkdir -p gpurun_out
timeout 600 python bench.py --config vqc20 --steps 3 --warmup 3 --no-aux > gpurun_out/bench_vqc20.log 2>&1; echo "vqc20 rc=$?"; tail -1 gpurun_out/bench_vqc20.log | cut -c1-1800
timeout 900 python bench.py --config vqc30 --steps 3 --warmup 3 --no-aux > gpurun_out/bench_vqc30.log 2>&1; echo "vqc30 rc=$?"; tail -1 gpurun_out/bench_vqc30.log | cut -c1-1800
