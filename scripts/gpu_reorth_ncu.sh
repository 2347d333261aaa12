#!/bin/bash
# Config C4 with the cooperative re-orthogonalisation kernel: timing, then the ncu launch list (durations only) and a
# section capture of k_reorth_chain.
OUT=gpurun_out/${1:-reorth_ncu}
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q -k "lanczos or reorth or quad" > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
tail -2 $OUT/pytest_gpu.log
CMD="python scripts/run_configs.py --only C4 --out $OUT/c4.jsonl"
$CMD > $OUT/plain.log 2>&1 || exit 1
tail -1 $OUT/c4.jsonl | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 24 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 600 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section LaunchStats --section Occupancy \
  --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_reorth_chain -s 60 -c 2 -o $OUT/prof_chain $CMD > $OUT/ncu_chain.log 2>&1
echo "ncu chain exit $?"
ncu -i $OUT/prof_chain.ncu-rep --page details > $OUT/ncu_chain_details.txt 2>&1
