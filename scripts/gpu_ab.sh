#!/bin/bash
# A/B of the step-A kernels and of the locality ordering on C2, plus the GPU tests (every command under its own timeout).
OUT=gpurun_out/${1:-ab}
mkdir -p $OUT
run() { echo "== $1"; shift; env "$@" timeout 300 python scripts/quick_perf.py $EXTRA 2>&1 | grep -v "rep 0" | tail -4; }
echo "== engine tests"; timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -5 | tee $OUT/pytest_engine.log
run "TMA + reorder(120)" SLQ_REORDER=1
run "TMA natural order" SLQ_REORDER=0
run "CSR + reorder(120)" SLQ_REORDER=1 SLQ_STEPA=csr
run "TMA + reorder(60)" SLQ_REORDER=1 SLQ_REORDER_TILE=60
run "TMA + reorder(240)" SLQ_REORDER=1 SLQ_REORDER_TILE=240
EXTRA="--k 64 --blocks 2" run "TMA + reorder k=64" SLQ_REORDER=1
EXTRA="--k 16" run "TMA + reorder k=16" SLQ_REORDER=1
EXTRA="--dtype float32" run "TMA + reorder fp32" SLQ_REORDER=1
