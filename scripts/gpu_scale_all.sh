#!/bin/bash
# 1 x 8-GPU box: the contract bench at N = 2, 4, 8 (C4 point-sharded, strong) and the C4 / C5 LM-rate lines at N = 4 / 8.
mkdir -p gpurun_out
run() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) "$@"; }
for n in 2 4 8; do run $n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; done
for n in 4 8; do run $n scripts/scale_bench.py --mode C4 --iters 10 > gpurun_out/c4_n$n.json 2> gpurun_out/c4_n$n.err; done
run 8 scripts/scale_bench.py --mode C5 --iters 10 > gpurun_out/c5_n8.json 2> gpurun_out/c5_n8.err
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/scale_gpus.csv
