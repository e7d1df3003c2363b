#!/bin/bash
# usage: gpu_scale2.sh "<N list>"  : scale_bench (C3, C4) and bench.py at every N
mkdir -p gpurun_out
: > gpurun_out/scale.jsonl; : > gpurun_out/bench_scale.jsonl
for n in $1; do
  for args in "--mode C3 --windows 4096" "--mode C4 --scale 1.0"; do
    if [ "$n" = "1" ]; then timeout 900 python scripts/scale_bench.py $args >> gpurun_out/scale.jsonl 2>> gpurun_out/scale.err
    else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 scripts/scale_bench.py $args >> gpurun_out/scale.jsonl 2>> gpurun_out/scale.err; fi
    echo "n=$n $args rc=$?"
  done
  if [ "$n" = "1" ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu >> gpurun_out/bench_scale.jsonl 2>> gpurun_out/scale.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $n --steps 20 --warmup 3 --no-cpu >> gpurun_out/bench_scale.jsonl 2>> gpurun_out/scale.err; fi
  echo "n=$n bench rc=$?"
done
cat gpurun_out/scale.jsonl | cut -c1-300; cat gpurun_out/bench_scale.jsonl | cut -c1-200; tail -5 gpurun_out/scale.err
