#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m qfedx_b200.benchmarks tiles > gpurun_out/tiles.log 2>&1; echo "tiles rc=$?"; python -c "
import json
for r in json.load(open('gpurun_out/tiles.log')): print(r['n_qubits'], r['batch'], r['tile_bits'], r['low_bits'], r['passes'], round(r['ms'],3), f\"{r['gates_per_s']:.3e}\", round(r['effective_gbs']), round(r['expz_ms'],3))
" || tail -20 gpurun_out/tiles.log
timeout 600 python -m qfedx_b200.benchmarks > gpurun_out/aux.log 2>&1; echo "aux rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/aux.log')); print(json.dumps(d['unitary_gemm_8q'])); print(json.dumps(d['sv20_pass_engine']))" || tail -20 gpurun_out/aux.log
