#!/bin/bash
# C3 (R-MAT Laplacian) A/B: row-balancing order on/off, for libslq.so and for any variants_*.so builds.
OUT=gpurun_out/${1:-c3ab}
mkdir -p $OUT
SCALE=${2:-24}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/pytest.log
for so in primate_b200/libslq.so primate_b200/variants_*.so; do
  [ -f $so ] || continue
  for r in ${REORDERS:-balance none}; do
    tag=$(basename $so .so)_$r
    echo "== $tag"
    SLQ_LIB=$PWD/$so timeout 900 python scripts/run_configs.py --only C3 --c3-scale $SCALE --c3-probes 64 --c3-reorder $r --out $OUT/$tag.json 2>&1 | tail -4
  done
done
