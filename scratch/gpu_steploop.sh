#!/bin/bash
mkdir -p gpurun_out
SBA_DEBUG_LS=2 SBA_TIMING=1 timeout 200 python scratch/diag.py C2 > gpurun_out/steploop.log 2>&1; echo rc=$?
grep -c "\[step\]" gpurun_out/steploop.log; grep "\[step\]" gpurun_out/steploop.log | tail -22; tail -3 gpurun_out/steploop.log
