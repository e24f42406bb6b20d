#!/bin/bash
# Experiments on the tiled kernel: PVA_TILED_DEBUG bits 1=no stores 2=no loads 4=no compute;
# PVA_TILED_RPS / PVA_TILED_OUTBUFS / PVA_TILED_STAGES force a tile configuration.
# CFGS="rps,bufs,stages ..."  DBG="0 3 ..."
run() { timeout 90 python bench.py $BENCH_ARGS --steps 2000 --warmup 30 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step']*1000,1),'us', d['config']['launch']['rows_per_tile'], d['config']['launch']['smem_bytes'])"; }
for cfg in ${CFGS:-4,3,2 4,2,2}; do IFS=, read rps bufs stages <<< "$cfg"
  for d in ${DBG:-0 1 2 4 3 5 6}; do PVA_TILED_RPS=$rps PVA_TILED_OUTBUFS=$bufs PVA_TILED_STAGES=$stages PVA_TILED_DEBUG=$d run "rps=$rps bufs=$bufs stages=$stages debug=$d"; done; done
