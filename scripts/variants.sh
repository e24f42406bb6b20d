#!/bin/bash
# A/B of launch-level switches on the default workload (graph and plain launches)
run() { python bench.py $2 --steps 3000 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 $2', round(d['ms_per_step']*1000,1),'us')"; }
for g in "--graph" ""; do
PVA_TILED_PDL=0 run "pdl=0" "$g"
PVA_TILED_PDL=1 run "pdl=1" "$g"
PVA_TILED_L2HINT=1 run "pdl=1 hint=1" "$g"
PVA_TILED_L2HINT=2 run "pdl=1 hint=2" "$g"
PVA_TILED_L2HINT=3 run "pdl=1 hint=3" "$g"
PVA_TILED_PDL=0 run "pdl=0 again" "$g"
done
