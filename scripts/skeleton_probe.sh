#!/bin/bash
# fixed cost of a launch: time the kernel with loads/compute/stores disabled
for b in ${BATCHES:-8}; do for d in ${DBG:-7 15 23 0 8}; do PVA_TILED_DEBUG=$d python bench.py --graph --batch $b --steps 2000 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('batch=$b debug=$d', round(d['ms_per_step']*1000,1),'us tiles', d['config']['launch']['tiles'], 'grid', d['config']['launch']['grid'])"; done; done
