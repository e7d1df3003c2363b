#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_chol_cluster' -s 3 -c 1 -f -o gpurun_out/full_chol python scratch/diag.py C2 > gpurun_out/ncu_chol.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_chol.log
