#!/bin/bash
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi2.txt
timeout 600 python -m pytest tests/test_parallel.py tests/test_shim.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -30 gpurun_out/pytest_gpu2.log
timeout 300 python scratch/diag.py C1 C2 > gpurun_out/diag.log 2>&1; cat gpurun_out/diag.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?"
cat gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
