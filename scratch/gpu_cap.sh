#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scratch/cap_sweep.py 96 112 128 144 160 176 192 208 224 240 > gpurun_out/cap_sweep.log 2>&1
cat gpurun_out/cap_sweep.log
