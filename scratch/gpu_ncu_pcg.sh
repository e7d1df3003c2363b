#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pcg_window' -s 6 -c 1 -f -o gpurun_out/full_pcg_c2 python scratch/diag.py C2 > gpurun_out/ncu_pcg.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pcg_window' -s 6 -c 1 -f -o gpurun_out/full_pcg_c1 python scratch/diag.py C1 >> gpurun_out/ncu_pcg.log 2>&1; echo "ncu rc=$?"
for c in c1 c2; do
ncu -i gpurun_out/full_pcg_$c.ncu-rep --page raw --csv > gpurun_out/pcg_raw_$c.csv 2>/dev/null
ncu -i gpurun_out/full_pcg_$c.ncu-rep --page source --csv > gpurun_out/pcg_source_$c.csv 2>/dev/null
done
SBA_TIMING=1 python scratch/diag.py C2 2>&1 | tail -8
