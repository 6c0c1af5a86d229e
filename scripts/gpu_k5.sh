#!/bin/bash
# K5 pass engine: numerics tests, throughput at 20 qubits x 16 states, one ncu capture with source counters.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_large.py -x -q --timeout 300 > gpurun_out/pytest_large.log 2>&1; echo "large tests rc=$?"; tail -3 gpurun_out/pytest_large.log | cut -c1-300
timeout 300 python - > gpurun_out/k5_bench.log 2>&1 <<'PY'
import json
from qfedx_b200.benchmarks import large_state_throughput as lt
print(json.dumps([lt(20, 2, 16), lt(20, 2, 64), lt(24, 2, 4), lt(27, 1, 1)], indent=1))
PY
echo "k5 bench rc=$?"; grep '"ms"\|n_qubits\|effective' gpurun_out/k5_bench.log
if [ "${NCU:-1}" = "1" ]; then
timeout 300 python scripts/profile_kernels.py large > gpurun_out/plain_prof_large.log 2>&1 && \
timeout 800 ncu --set full --clock-control none --import-source on -k regex:'sv_large_pass' -c 8 -o /tmp/prof_k5 python scripts/profile_kernels.py large > gpurun_out/ncu_k5.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_k5.log | cut -c1-200
ncu -i /tmp/prof_k5.ncu-rep --page raw --csv > gpurun_out/prof_k5_raw.csv 2>/dev/null
ncu -i /tmp/prof_k5.ncu-rep --page source --csv --kernel-name regex:sv_large_pass_v2 > gpurun_out/prof_k5_source.csv 2>/dev/null; ls -la gpurun_out/prof_k5_*.csv
fi
