#!/bin/bash
# ncu --set full capture of the chunk kernels on the bench workload (C2), after the same command has exited 0 without ncu.
# Usage (under gpurun): bash scripts/gpu_ncu_chunks.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
export SBA_NO_GRAPH=1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > gpurun_out/ncu_plain_${tag}.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_fused_persist|k_fused_chunk|k_lin_chunk|k_schur_chunk' -s 4 -c 12 -f -o gpurun_out/chunks_${tag} $CMD > gpurun_out/ncu_${tag}.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_${tag}.log
