#!/bin/bash
# A/B of tile configurations on the default workload
run() { timeout 90 python bench.py --steps 3000 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step']*1000,1),'us', d['config']['launch']['rows_per_tile'], d['config']['launch']['smem_bytes'])"; }
PVA_TILED_RPS=4 PVA_TILED_RPST=4 PVA_TILED_OUTBUFS=3 run "rps4 rpst4 bufs3"
PVA_TILED_RPS=8 PVA_TILED_RPST=4 run "rps8 rpst4"
PVA_TILED_RPS=8 PVA_TILED_RPST=8 run "rps8 rpst8"
PVA_TILED_RPS=8 PVA_TILED_RPST=2 run "rps8 rpst2"
PVA_TILED_RPS=8 PVA_TILED_RPST=4 PVA_TILED_DEBUG=3 run "rps8 rpst4 compute only"
PVA_TILED_RPS=4 PVA_TILED_RPST=4 PVA_TILED_OUTBUFS=3 PVA_TILED_DEBUG=3 run "rps4 compute only"
run "default"
