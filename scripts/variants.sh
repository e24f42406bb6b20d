#!/bin/bash
# A/B of launch-level switches on the default workload
run() { timeout 90 python bench.py $2 --steps ${STEPS:-3000} --warmup 40 --no-cpu-baseline --e2e-steps 5 2>gpurun_out/variant.err | python -c "
import json,sys
s=sys.stdin.read()
try:
    d=json.loads(s); print('$1 $2', round(d['ms_per_step']*1000,2),'us', 'overlapped', round(d['overlapped_launches']['ms_per_step']*1000,2))
except Exception: print('$1 FAILED:', open('gpurun_out/variant.err').read().strip().splitlines()[-1][:150] if s=='' else s[:100])"; }
for g in "" "--no-graph"; do
PVA_TILED_PDL=2 run "pdl wait at start" "$g"
PVA_TILED_PDL=1 run "pdl wait before first store" "$g"
PVA_TILED_PDL=0 run "no pdl" "$g"
done
