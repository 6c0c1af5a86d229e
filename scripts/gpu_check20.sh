#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -c "
import json, torch
from qfedx_b200.benchmarks import large_state_throughput
for reg in (True, False):
    for n,L,b in ((20,2,16),(20,2,4),(24,1,1)):
        r=large_state_throughput(n,L,b,register_engine=reg); print(reg, n, b, r['passes'], round(r['ms'],3), f\"{r['gates_per_s']:.3e}\", round(r['effective_gbs']))
"
