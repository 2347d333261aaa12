#!/bin/bash
# ncu of the step-A kernel on the R-MAT Laplacian of config C3 (scale $2, default 24), row-balancing order on.
OUT=gpurun_out/${1:-ncuc3}
SCALE=${2:-24}
mkdir -p $OUT
CMD="python scripts/run_configs.py --only C3 --c3-scale $SCALE --c3-probes 32 --c3-reorder ${3:-balance} --out $OUT/c3.jsonl"
$CMD > $OUT/plain.log 2>&1 &&
timeout 1500 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section MemoryWorkloadAnalysis_Tables --section WarpStateStats \
  --section SchedulerStats --section LaunchStats --section Occupancy --section SourceCounters \
  --metrics sm__cycles_active.min,sm__cycles_active.max,sm__cycles_active.avg,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct \
  --clock-control none --import-source on -k regex:'k_stepA_tma' -s 40 -c 1 -o $OUT/prof_c3 $CMD > $OUT/ncu.log 2>&1
echo "exit $?"; tail -2 $OUT/plain.log | cut -c1-600
