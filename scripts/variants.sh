#!/bin/bash
# A/B of tile configurations on the default workload
run() { timeout 90 python bench.py --steps ${STEPS:-3000} --warmup 40 --no-cpu-baseline --e2e-steps 5 2>gpurun_out/variant.err | python -c "
import json,sys
s=sys.stdin.read()
try:
    d=json.loads(s); print('$1', round(d['ms_per_step']*1000,1),'us', d['config']['launch']['smem_bytes'])
except Exception: print('$1 FAILED:', open('gpurun_out/variant.err').read().strip().splitlines()[-1][:150] if s=='' else s[:100])"; }
PVA_TILED_GROUPS=1 run "1 group rps4"
PVA_TILED_GROUPS=2 run "2 groups rps8"
for d in 3 4 6 1; do PVA_TILED_GROUPS=2 PVA_TILED_DEBUG=$d run "2 groups debug=$d"; done
PVA_TILED_GROUPS=2 PVA_TILED_PDL=0 run "2 groups nopdl"
PVA_TILED_GROUPS=2 PVA_TILED_L2HINT=0 run "2 groups nohint"
