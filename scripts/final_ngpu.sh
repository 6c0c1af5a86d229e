# usage: bash scripts/final_ngpu.sh N   (inside a gpurun --gpus N call): the round's multi-GPU measurements
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "8" ]; then
  $TR --master-port 29701 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/f${N}_cfed.json 2> gpurun_out/f${N}_cfed.err; tail -1 gpurun_out/f${N}_cfed.json | cut -c1-200
  for c in vqc8 vqc20 vqc30; do $TR --master-port 29702 bench.py --gpus $N --config $c --steps 5 --warmup 3 --no-aux > gpurun_out/f${N}_$c.json 2> gpurun_out/f${N}_$c.err; tail -1 gpurun_out/f${N}_$c.json | cut -c1-160; done
  timeout 600 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/f${N}_multigpu.log 2>&1; tail -2 gpurun_out/f${N}_multigpu.log
else
  $TR --master-port 29701 bench.py --gpus $N --steps 10 --warmup 3 --no-aux > gpurun_out/f${N}_cfed.json 2> gpurun_out/f${N}_cfed.err; tail -1 gpurun_out/f${N}_cfed.json | cut -c1-200
  $TR --master-port 29702 bench.py --gpus $N --config vqc8 --steps 10 --warmup 3 --no-aux > gpurun_out/f${N}_vqc8.json 2> gpurun_out/f${N}_vqc8.err; tail -1 gpurun_out/f${N}_vqc8.json | cut -c1-160
fi
