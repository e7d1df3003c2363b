#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_step_loop' -s 4 -c 1 -f -o gpurun_out/full_step python scratch/diag.py C2 > gpurun_out/ncu_step.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_step.log
