#!/bin/bash
# Source-level ncu capture of one TinyCNN kernel (default fc1_bwd_train) inside the graph-replayed round.
mkdir -p gpurun_out
K=${1:-fc1_bwd_train}
export QFX_BENCH_SCALE=0.05
timeout 300 python bench.py --steps 2 --warmup 3 --no-aux > gpurun_out/plain_cnn.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:"$K" --launch-skip 40 -c 2 -o /tmp/prof_src python bench.py --steps 2 --warmup 3 --no-aux > gpurun_out/ncu_src.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_src.log | cut -c1-200
ncu -i /tmp/prof_src.ncu-rep --page source --csv > gpurun_out/prof_src_$K.csv 2>/dev/null; ls -la gpurun_out/prof_src_$K.csv
