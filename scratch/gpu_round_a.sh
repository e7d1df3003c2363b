#!/bin/bash
# tests + bench + ncu launch list + one full capture of the two chunk kernels
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_lin_chunk|k_schur_chunk|k_assemble_schur|k_assemble_cams|k_chol_cluster|k_step_loop' -s 12 -c 6 -f -o gpurun_out/full_chunks python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out
timeout 600 python scripts/scale_bench.py --mode C3 --windows 4096 > gpurun_out/scale_c3_n1.json 2> gpurun_out/scale_c3.err; echo "c3 rc=$?"; cat gpurun_out/scale_c3_n1.json
timeout 600 python scripts/scale_bench.py --mode C4 --scale 1.0 > gpurun_out/scale_c4_n1.json 2> gpurun_out/scale_c4.err; echo "c4 rc=$?"; cat gpurun_out/scale_c4_n1.json
timeout 300 python scratch/diag.py C1 C2 > gpurun_out/diag.log 2>&1; cat gpurun_out/diag.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
