#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --config vqc8 --steps 2 --warmup 1 --no-aux > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vqc_local_train -c 1 -o gpurun_out/prof_vqc_train_latency python bench.py --config vqc8 --steps 2 --warmup 1 --no-aux > gpurun_out/ncu_train.log 2>&1; echo "ncu rc=$?"
QFX_BENCH_CLIENTS=1 timeout 300 python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/plain2.log 2>&1 && \
QFX_BENCH_CLIENTS=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 45 --csv --log-file gpurun_out/launches_cfed_k1.csv python bench.py --config cfed --steps 1 --warmup 1 --no-aux > gpurun_out/ncu_cfed_k1.log 2>&1; echo "ncu2 rc=$?"
tail -1 gpurun_out/plain2.log | cut -c1-200
