#!/bin/bash
# Round profile: bench lines for the three workloads, ncu launch list and one --set full capture
# of the headline kernel.  Outputs land in gpurun_out/ (copied into profiles/ afterwards).
set -x
for w in slowfast_train_b8 slowfast_val_b8 single_pathway_b64; do
  python bench.py --workload $w --steps 3000 --warmup 40 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
CMD="python bench.py --steps 24 --warmup 8 --no-cpu-baseline --e2e-steps 5 --no-graph"
$CMD > gpurun_out/plain_launches.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pva_tiled -s 12 -c 2 -f -o gpurun_out/prof_r1_tiled $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
