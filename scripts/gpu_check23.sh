#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_statevector.py tests/test_gpu_unitary_gemm.py tests/test_gpu_large.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python -c "
import json
from qfedx_b200.benchmarks import small_circuit_throughput
print(json.dumps(small_circuit_throughput(8,6,1<<15)))
print(json.dumps(small_circuit_throughput(4,2,1<<18)))
print(json.dumps(small_circuit_throughput(10,4,1<<13)))
"
