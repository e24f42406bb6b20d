#!/bin/bash
# BASELINE.json configs[3]: single-pathway transform (16 frames, crop 224), bandwidth sweep over
# batch and short-side size.  CSV on stdout.
echo "batch,size,us_per_launch,clips_per_s,algorithmic_GBps,frac_of_measured_peak"
for b in 8 16 32 64 128; do for s in 256 288 320; do
python bench.py --workload single_pathway_b64 --batch $b --size $s --steps 1500 --warmup 40 --no-cpu-baseline --e2e-steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$b,$s,%.1f,%.0f,%.0f,%.3f' % (d['ms_per_step']*1000, d['value'], r['achieved'], r['frac']))"
done; done
