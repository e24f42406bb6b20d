#!/bin/bash
# loads-only / loads+stores ablation (PVA_TILED_DEBUG) for column-cropped vs linear loads
mkdir -p gpurun_out
out=gpurun_out/ab_colcrop_ablate.txt
: > $out
for dbg in 0 4 5 3; do
  for cc in 1 0; do
    PVA_TILED_DEBUG=$dbg PVA_TILED_COLCROP=$cc timeout 120 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline --e2e-steps 0 \
      2>> gpurun_out/ab_colcrop.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('debug $dbg colcrop $cc us/launch', round(1000*d['ms_per_step'],2), 'overlapped us', round(1000*d['overlapped_launches']['ms_per_step'],2))
" >> $out 2>&1
  done
done
cat $out
