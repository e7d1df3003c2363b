#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/stepgrid.log
for g in 592 444 296 148; do
  echo "SBA_STEP_GRID=$g" >> gpurun_out/stepgrid.log
  SBA_STEP_GRID=$g timeout 100 python scratch/diag.py C2 2>&1 | tail -2 >> gpurun_out/stepgrid.log
done
cat gpurun_out/stepgrid.log
