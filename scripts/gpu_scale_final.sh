#!/bin/bash
# 1 x 8-GPU box: the contract bench at N = 4, 8 and the C4 / C5 LM-rate lines at N = 8 (N = 2 runs on a 2-GPU box).
mkdir -p gpurun_out
run() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) "$@"; }
for n in 4 8; do run $n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; done
run 8 scripts/scale_bench.py --mode C4 --iters 10 > gpurun_out/c4_n8.json 2> gpurun_out/c4_n8.err
run 8 scripts/scale_bench.py --mode C5 --iters 10 > gpurun_out/c5_n8.json 2> gpurun_out/c5_n8.err
