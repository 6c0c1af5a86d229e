#!/bin/bash
# K7 check: numerics tests, latency at model sizes and at the 10^8-float roofline size, then one ncu capture.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train_agg.py tests/test_zz_no_alloc.py -x -q --timeout 200 > gpurun_out/pytest_agg.log 2>&1; echo "agg tests rc=$?"; tail -3 gpurun_out/pytest_agg.log | cut -c1-300
timeout 300 python - > gpurun_out/agg_bench.log 2>&1 <<'PY'
import json
from qfedx_b200.benchmarks import aggregation_latency as al
res = {"vqc": al(64, 102), "tinycnn_dp_secagg": al(8, 113859, True), "tinycnn_dp": al(8, 113859, False),
       "1e8_dp": al(4, 25_000_000, iters=10), "1e8_clip": al(4, 25_000_000, sigma=0.0, iters=10),
       "1e8_weights": al(4, 25_000_000, mode="weights", clip=0.0, sigma=0.0, iters=10),
       "1e8_k16_weights": al(16, 6_250_000, mode="weights", clip=0.0, sigma=0.0, iters=10)}
print(json.dumps(res, indent=1))
PY
echo "agg bench rc=$?"; grep '"ms"\|^ "' gpurun_out/agg_bench.log
timeout 300 python scripts/profile_kernels.py agg > gpurun_out/plain_prof_agg.log 2>&1 && \
timeout 800 ncu --set full --clock-control none --import-source on -k regex:'fedavg_fused_kernel' -c 12 -o /tmp/prof_agg python scripts/profile_kernels.py agg > gpurun_out/ncu_agg.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_agg.log | cut -c1-200
ncu -i /tmp/prof_agg.ncu-rep --page raw --csv > gpurun_out/prof_agg_raw.csv 2>/dev/null; ls -la /tmp/prof_agg.ncu-rep gpurun_out/prof_agg_raw.csv
