#!/bin/bash
# Tuning aid: run scripts/quick_perf.py against every libslq build variant (primate_b200/variants_*.so), with the
# locality ordering on and off.
for so in primate_b200/libslq.so primate_b200/variants_*.so; do
  for r in 1 0; do
    echo "== $so reorder=$r"
    SLQ_REORDER=$r SLQ_LIB=$PWD/$so timeout 300 python scripts/quick_perf.py "$@" 2>&1 | tail -5 | grep -v "rep 0" | grep -v "rep 1" | tail -2
  done
done
