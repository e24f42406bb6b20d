#!/bin/bash
# A/B of environment switches on the headline workload.  usage: ab_env.sh "VAR=.. VAR=.." ...   (optional BENCH_ARGS)
mkdir -p gpurun_out
out=gpurun_out/ab_env.txt
: > $out
for spec in "$@"; do
  env $spec timeout 120 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline --e2e-steps 0 $BENCH_ARGS \
    2>> gpurun_out/ab_env.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$spec', 'us/launch', round(1000*d['ms_per_step'],2), 'overlapped us', round(1000*d['overlapped_launches']['ms_per_step'],2))
" >> $out 2>&1
done
cat $out
