#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python scratch/diag.py C1 C2 > gpurun_out/diag.log 2>&1; cat gpurun_out/diag.log
SBA_NO_DIRECT=1 timeout 300 python scratch/diag.py C1 C2 > gpurun_out/diag_pcg.log 2>&1; cat gpurun_out/diag_pcg.log
