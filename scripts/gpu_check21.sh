#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/profile_kernels.py large > gpurun_out/plain_l.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k regex:sv_large_pass -s 1 -c 2 -o /tmp/prof_l python scripts/profile_kernels.py large > gpurun_out/ncu_l.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/prof_l.ncu-rep --page raw --csv > gpurun_out/prof_large_v2_raw.csv 2>/dev/null; ls -la gpurun_out/prof_large_v2_raw.csv
