#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_lin_chunk|k_schur_chunk' -s 4 -c 2 -f -o gpurun_out/full_chunks python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/full_chunks.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2>/dev/null
ncu -i gpurun_out/full_chunks.ncu-rep --page source --csv > gpurun_out/full_source.csv 2>/dev/null
ls -la gpurun_out | head -30
