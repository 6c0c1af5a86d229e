#!/bin/bash
mkdir -p gpurun_out
for sh in array feistel array feistel; do
QFX_SHUFFLE=$sh timeout 600 python bench.py --config cfed --steps 5 --warmup 3 --no-aux > gpurun_out/bench_cfed_$sh.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench_cfed_$sh.log').read().strip().splitlines()[-1]); print('$sh', d['ms_per_step'], d['value'], d['e2e']['value'])"
done
