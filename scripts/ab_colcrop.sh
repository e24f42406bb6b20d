#!/bin/bash
# A/B of the column-cropped tensor-map loads (PVA_TILED_COLCROP=1, default) against the linear band
# copies (=0): kernel-only bench lines for the three workloads.  Output: gpurun_out/ab_colcrop.txt
mkdir -p gpurun_out
out=gpurun_out/ab_colcrop.txt
: > $out
for wl in slowfast_train_b8 slowfast_val_b8 single_pathway_b64; do
  for cc in 1 0; do
    echo "== $wl COLCROP=$cc" >> $out
    PVA_TILED_COLCROP=$cc timeout 120 python bench.py --workload $wl --steps 3000 --warmup 100 \
      --no-cpu-baseline --e2e-steps 0 2>> gpurun_out/ab_colcrop.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('us/launch', round(1000*d['ms_per_step'],2), 'frac', round(d['roofline']['frac'],4), 'clips/s', round(d['value']), 'overlapped us', round(1000*d['overlapped_launches']['ms_per_step'],2), d['config']['launch'])
" >> $out 2>&1
  done
done
cat $out
