#!/bin/bash
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none --import-source on -k regex:'k_chol_cluster|k_step_loop' -s 6 -c 2 -f -o gpurun_out/full_e2e python scratch/diag.py C2 > gpurun_out/ncu_e2e.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/full_e2e.ncu-rep --page raw --csv > gpurun_out/full_e2e_raw.csv 2>/dev/null
rm -f gpurun_out/full_e2e.ncu-rep
tail -3 gpurun_out/ncu_e2e.log; wc -c gpurun_out/full_e2e_raw.csv
