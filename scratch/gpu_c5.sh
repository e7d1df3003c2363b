#!/bin/bash
# usage: gpu_c5.sh N [scale]
mkdir -p gpurun_out
n=$1; sc=${2:-1.0}
if [ "$n" = "1" ]; then timeout ${TMO:-420} python scripts/scale_bench.py --mode C5 --scale $sc --steps 5 --iters ${3:-5} > gpurun_out/c5_n$n.json 2> gpurun_out/c5_n$n.err
else timeout ${TMO:-420} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 scripts/scale_bench.py --mode C5 --scale $sc --steps 5 --iters ${3:-5} > gpurun_out/c5_n$n.json 2> gpurun_out/c5_n$n.err; fi
echo "rc=$?"; cat gpurun_out/c5_n$n.json; tail -5 gpurun_out/c5_n$n.err; nvidia-smi --query-gpu=memory.used --format=csv | head -3
