#!/bin/bash
# time per launch vs batch size (headline transform): the intercept is the fixed cost of a launch
for b in 2 4 8 16 32; do timeout 120 python bench.py --batch $b --steps 2000 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('batch=$b', round(d['ms_per_step']*1000,1),'us', 'tiles', d['impl_detail']['launch']['tiles'], 'frac', round(d['roofline']['frac'],3))"; done
