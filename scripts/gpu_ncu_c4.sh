#!/bin/bash
# ncu of the dense-operator kernels (config C4: 32 768^2 Gram matrix, full re-orthogonalisation): k_dense_tma and the
# re-orthogonalisation kernel.
OUT=gpurun_out/${1:-ncuc4}
mkdir -p $OUT
CMD="python scripts/run_configs.py --only C4 --out $OUT/c4.jsonl"
$CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section SchedulerStats \
  --section LaunchStats --section Occupancy --section ComputeWorkloadAnalysis \
  --metrics dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fp64.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum \
  --clock-control none --import-source on -s 240 -c 8 -o $OUT/prof_c4 $CMD > $OUT/ncu.log 2>&1
echo "exit $?"; tail -1 $OUT/plain.log | cut -c1-400
