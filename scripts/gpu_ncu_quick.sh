#!/bin/bash
# ncu --set full of the step-A kernel on C2 through scripts/quick_perf.py (1 block), natural and locality order.
OUT=gpurun_out/${1:-ncuq}
mkdir -p $OUT
CMD="python scripts/quick_perf.py --blocks 1"
SLQ_REORDER=0 $CMD > $OUT/plain_nat.log 2>&1 &&
SLQ_REORDER=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_stepA_tma' -s 120 -c 2 -o $OUT/prof_tma_nat $CMD > $OUT/ncu_nat.log 2>&1
echo "nat exit $?"
SLQ_REORDER=1 $CMD > $OUT/plain_reo.log 2>&1 &&
SLQ_REORDER=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_stepA_tma' -s 120 -c 2 -o $OUT/prof_tma_reo $CMD > $OUT/ncu_reo.log 2>&1
echo "reo exit $?"
ls -la $OUT
