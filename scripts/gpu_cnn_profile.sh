#!/bin/bash
# ncu capture of one TinyCNN training step (8 kernels, all 8 clients batched) inside the graph-replayed round.
mkdir -p gpurun_out
export QFX_BENCH_SCALE=0.05
timeout 300 python bench.py --steps 2 --warmup 3 --no-aux > gpurun_out/plain_cnn.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:'_train|_umma' --launch-skip 160 -c 16 -o /tmp/prof_cnn python bench.py --steps 2 --warmup 3 --no-aux > gpurun_out/ncu_cnn.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_cnn.log | cut -c1-200
ncu -i /tmp/prof_cnn.ncu-rep --page raw --csv > gpurun_out/prof_cnn_raw.csv 2>/dev/null; ls -la gpurun_out/prof_cnn_raw.csv
