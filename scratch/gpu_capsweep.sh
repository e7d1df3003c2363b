#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/capsweep.log
for cap in 240 184 176 168 160; do
  SBA_OBS_CAP=$cap timeout 100 python bench.py --steps 20 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('cap $cap value %.3e us/step %.2f' % (b['value'], 1e3*b['ms_per_step']), b['stages_us'], 'e2e ms %.2f' % b['e2e']['ms_per_solve'], 'cost', b['e2e']['final_cost'], 'it', b['e2e']['lm_iterations'])" | tee -a gpurun_out/capsweep.log
done
