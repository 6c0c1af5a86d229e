#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_large.py -m gpu -x -q --timeout 120 > gpurun_out/pytest_large.log 2>&1; echo "pytest large rc=$?"; tail -15 gpurun_out/pytest_large.log | cut -c1-300
timeout 300 python -c "
import json
from qfedx_b200.benchmarks import large_state_throughput
print(json.dumps(large_state_throughput(20,2,16)))
print(json.dumps(large_state_throughput(26,1,1)))
"
