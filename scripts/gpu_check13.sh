#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_unitary_gemm.py -m gpu -x -q > gpurun_out/pytest_k2.log 2>&1; echo "pytest k2 rc=$?"; tail -5 gpurun_out/pytest_k2.log | cut -c1-300
timeout 300 python -c "
import json
from qfedx_b200.benchmarks import unitary_gemm_throughput
print(json.dumps(unitary_gemm_throughput(8,6,1<<15)))
print(json.dumps(unitary_gemm_throughput(10,4,1<<13)))
"
