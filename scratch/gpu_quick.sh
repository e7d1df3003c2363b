#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    b=json.load(open('gpurun_out/bench.json'))
    print('value %.3e ms/step %.4f e2e %.3e ms/solve %.2f it/s %.1f' % (b['value'], b['ms_per_step'], b['e2e']['value'], b['e2e']['ms_per_solve'], b['lm_iters_per_s']))
    print('stages', b['stages_us']); print('e2e dev', b['e2e']['device_ms'], b['e2e']['ms_breakdown'])
    print('roofline', b['roofline'])
except Exception as e:
    print('bench parse failed', e); print(open('gpurun_out/bench.err').read()[-2000:])
PY
timeout 300 python scratch/diag.py C1 C2 > gpurun_out/diag.log 2>&1; cat gpurun_out/diag.log
