#!/bin/bash
# Dense kernel A/B on config C4: CTA heights 128 / 112 / 96 rows (SLQ_DENSE_RPT) after the GPU parity tests.
OUT=gpurun_out/${1:-dense_ab}
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/gpu.csv 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
for rpt in auto 8 7 6; do
  if [ $rpt = auto ]; then unset SLQ_DENSE_RPT; else export SLQ_DENSE_RPT=$rpt; fi
  timeout 600 python scripts/run_configs.py --only C4 --out $OUT/c4_rpt$rpt.jsonl > $OUT/c4_rpt$rpt.log 2>&1
  echo "rpt $rpt exit $?"; tail -1 $OUT/c4_rpt$rpt.jsonl | cut -c1-330
done
