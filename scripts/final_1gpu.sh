mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/f1_pytest.log 2>&1; tail -2 gpurun_out/f1_pytest.log
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/f1_ref.json 2> gpurun_out/f1_ref.err; tail -1 gpurun_out/f1_ref.json | cut -c1-200
python bench.py --steps 10 --warmup 3 > gpurun_out/f1_cfed.json 2> gpurun_out/f1_cfed.err; tail -1 gpurun_out/f1_cfed.json | cut -c1-200
for c in vqc8 vqc4 vqc20; do python bench.py --config $c --steps 10 --warmup 3 --no-aux > gpurun_out/f1_$c.json 2> gpurun_out/f1_$c.err; tail -1 gpurun_out/f1_$c.json | cut -c1-160; done
python bench.py --config vqc20 --grad adjoint --steps 10 --warmup 3 --no-aux > gpurun_out/f1_vqc20_adj.json 2> gpurun_out/f1_vqc20_adj.err; tail -1 gpurun_out/f1_vqc20_adj.json | cut -c1-160
python scripts/profile_vqc_train.py 8 8 > gpurun_out/f1_vqcprof8.txt 2>&1
python scripts/profile_vqc_train.py 4 > gpurun_out/f1_vqcprof4_all.txt 2>&1
