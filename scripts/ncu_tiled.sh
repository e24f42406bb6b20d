#!/bin/bash
# ncu --set full capture of the tiled kernel (optionally under a PVA_TILED_DEBUG ablation).
# usage: scripts/ncu_tiled.sh <out-name> ; env PVA_TILED_* are passed through
set -e
CMD="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 5"
$CMD > gpurun_out/plain_$1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pva_tiled -s 10 -c 2 -f -o gpurun_out/$1 $CMD > gpurun_out/ncu_$1.log 2>&1
tail -2 gpurun_out/ncu_$1.log
