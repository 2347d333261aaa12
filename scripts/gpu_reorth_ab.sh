#!/bin/bash
# A/B of the re-orthogonalisation paths on config C4 (dense 32 768^2, deg 100, orth 100): the cooperative chain kernel
# against the per-vector launches (SLQ_REORTH=launches), after the GPU parity tests.
OUT=gpurun_out/${1:-reorth}
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/gpu.csv 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
for mode in chain launches; do
  SLQ_REORTH=$mode timeout 600 python scripts/run_configs.py --only C4 --out $OUT/c4_$mode.jsonl > $OUT/c4_$mode.log 2>&1
  echo "$mode exit $?"; tail -1 $OUT/c4_$mode.jsonl | cut -c1-600
done
