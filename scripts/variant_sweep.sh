#!/bin/bash
# Tuning aid: run scripts/quick_perf.py against every libslq build variant (primate_b200/variants_*.so).
for so in primate_b200/libslq.so primate_b200/variants_*.so; do
  echo "== $so"
  SLQ_LIB=$PWD/$so timeout 300 python scripts/quick_perf.py "$@" 2>&1 | tail -5 | grep -v "rep 0"
done
