#!/bin/bash
# column-cropped loads vs linear band copies at several batch sizes (steady-state cost per clip vs
# fixed cost per launch).  Output: gpurun_out/ab_colcrop_batch.txt
mkdir -p gpurun_out
out=gpurun_out/ab_colcrop_batch.txt
: > $out
for b in 4 8 16 32; do
  for cc in 1 0; do
    PVA_TILED_COLCROP=$cc timeout 120 python bench.py --batch $b --steps 2000 --warmup 100 \
      --no-cpu-baseline --e2e-steps 0 2>> gpurun_out/ab_colcrop.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('batch $b colcrop $cc us/launch', round(1000*d['ms_per_step'],2), 'overlapped us', round(1000*d['overlapped_launches']['ms_per_step'],2))
" >> $out 2>&1
  done
done
cat $out
