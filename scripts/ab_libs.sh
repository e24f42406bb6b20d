#!/bin/bash
# same-box A/B of library builds (PVA_LIBRARY) on the headline workload.  usage: ab_libs.sh "<env> <lib>" ...
mkdir -p gpurun_out
out=gpurun_out/ab_libs.txt
: > $out
for spec in "$@"; do
  env_s=${spec%% *}; lib=${spec##* }
  for rep in 1 2; do
    env $env_s PVA_LIBRARY=$lib timeout 120 python bench.py --steps 3000 --warmup 100 --no-cpu-baseline --e2e-steps 0 \
      2>> gpurun_out/ab_libs.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$spec', 'us/launch', round(1000*d['ms_per_step'],2), 'overlapped us', round(1000*d['overlapped_launches']['ms_per_step'],2))
" >> $out 2>&1
  done
done
cat $out
