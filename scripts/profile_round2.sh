#!/bin/bash
# Round-2 profile: bench lines for the three workloads under the driver's protocol and in a long run,
# the ncu launch list, steady-state DRAM traffic per launch for the three workloads, and one
# `ncu --set full` capture of the headline kernel.  Outputs land in gpurun_out/ (summaries are copied
# into profiles/ afterwards).  Every ncu command runs only after the same command exited 0 without ncu.
set -x
OUT=gpurun_out
for w in slowfast_train_b8 slowfast_val_b8 single_pathway_b64; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $OUT/r2_bench_${w}_k20.json 2> $OUT/r2_bench_${w}_k20.err
  python bench.py --workload $w --steps 3000 --warmup 40 --no-cpu-baseline --e2e-steps 5 > $OUT/r2_bench_${w}_k3000.json 2> $OUT/r2_bench_${w}_k3000.err
done
# launch list (cold, serialised): the kernel's share of the step
CMD="python bench.py --steps 24 --warmup 8 --no-cpu-baseline --e2e-steps 5 --overlap-streams 0 --no-rehearsal --no-clocks"
$CMD > $OUT/r2_plain_launches.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2_launches.csv $CMD > $OUT/r2_ncu_launches.log 2>&1
# steady-state DRAM traffic: caches NOT flushed between launches, 12 consecutive launches on rotating buffers
M=dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum
for w in slowfast_train_b8 slowfast_val_b8 single_pathway_b64; do
  CMDW="python bench.py --workload $w --steps 24 --warmup 8 --no-cpu-baseline --e2e-steps 5 --overlap-streams 0 --no-rehearsal --no-clocks"
  $CMDW > $OUT/r2_plain_$w.log 2>&1 && \
  ncu --metrics $M --cache-control none --clock-control none -k regex:pva_tiled -s 12 -c 12 --csv --log-file $OUT/r2_traffic_$w.csv $CMDW > $OUT/r2_ncu_traffic_$w.log 2>&1
done
# one full capture of the headline kernel
$CMD > $OUT/r2_plain_full.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pva_tiled -s 14 -c 2 -f -o $OUT/r2_prof_tiled $CMD > $OUT/r2_ncu_full.log 2>&1
tail -2 $OUT/r2_ncu_full.log
# the same traffic capture with every launch on all SMs (no PVA_FLAG_OUTPUTS_DISJOINT): what ncu, which
# serialises launches, sees of a full-chip launch in isolation
CMDS="python bench.py --steps 24 --warmup 8 --no-cpu-baseline --e2e-steps 5 --overlap-streams 0 --no-rehearsal --no-clocks --serialize-stores"
$CMDS > $OUT/r2_plain_fullgrid.log 2>&1 && \
ncu --metrics $M --cache-control none --clock-control none -k regex:pva_tiled -s 12 -c 12 --csv --log-file $OUT/r2_traffic_fullgrid_slowfast_train_b8.csv $CMDS > $OUT/r2_ncu_traffic_fullgrid.log 2>&1
