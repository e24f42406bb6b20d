#!/bin/bash
# A/B of launch-level switches on the default workload
run() { python bench.py $2 --steps 3000 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 $2', round(d['ms_per_step']*1000,1),'us', d['config']['launch']['smem_bytes'])"; }
PVA_TILED_LOAD_ROWS=0 run "bulk-band loads"
PVA_TILED_LOAD_ROWS=1 run "row loads"
PVA_TILED_LOAD_ROWS=1 PVA_TILED_DEBUG=5 run "row loads, loads only"
PVA_TILED_LOAD_ROWS=1 PVA_TILED_DEBUG=4 run "row loads, no arithmetic"
PVA_TILED_LOAD_ROWS=0 PVA_TILED_DEBUG=4 run "bulk loads, no arithmetic"
PVA_TILED_LOAD_ROWS=1 PVA_TILED_STAGES=3 PVA_TILED_OUTBUFS=3 run "row loads 3 stages"
