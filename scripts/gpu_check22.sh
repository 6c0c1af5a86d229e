#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_large.py -m gpu -x -q --timeout 200 > gpurun_out/pytest_large.log 2>&1; echo "pytest large rc=$?"; tail -12 gpurun_out/pytest_large.log | cut -c1-300
timeout 600 python -m qfedx_b200 train --preset vqc20 --set data.synthetic_scale=0.004 --set fl.num_rounds=1 --set fl.quiet=true --set model.n_qubits=14 > gpurun_out/cli_vqc20.log 2>&1; echo "cli vqc20(14q) rc=$?"; tail -2 gpurun_out/cli_vqc20.log | cut -c1-300
timeout 600 python -m qfedx_b200 train --preset vqc8 --set data.synthetic_scale=0.05 --set fl.num_rounds=2 --set fl.quiet=true > gpurun_out/cli_vqc8.log 2>&1; echo "cli vqc8 rc=$?"; tail -1 gpurun_out/cli_vqc8.log | cut -c1-300
timeout 600 python -m qfedx_b200 train --preset cfed --set data.synthetic_scale=0.05 --set fl.num_rounds=2 --set fl.quiet=true > gpurun_out/cli_cfed.log 2>&1; echo "cli cfed rc=$?"; tail -1 gpurun_out/cli_cfed.log | cut -c1-300
bash scripts/sanitize.sh memcheck
