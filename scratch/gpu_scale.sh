#!/bin/bash
# usage: gpu_scale.sh "<N list>"   (box must have max(N) GPUs)
mkdir -p gpurun_out
: > gpurun_out/scale.jsonl
for n in $1; do
  for args in "--mode C3 --windows 4096" "--mode C4 --scale 1.0"; do
    if [ "$n" = "1" ]; then timeout 900 python scripts/scale_bench.py $args >> gpurun_out/scale.jsonl 2>> gpurun_out/scale.err
    else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 scripts/scale_bench.py $args >> gpurun_out/scale.jsonl 2>> gpurun_out/scale.err; fi
    echo "n=$n $args rc=$?"
  done
done
cat gpurun_out/scale.jsonl; tail -5 gpurun_out/scale.err
