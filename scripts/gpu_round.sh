#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the ncu launch list and one --set full capture of the top kernels.
# Usage (from the repo root, on the GPU box):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,power.limit --format=csv > $OUT/gpu.csv 2>&1
echo "== pytest -m gpu"
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
echo "== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" | tee -a $OUT/smoke.log
tail -3 $OUT/smoke.log
echo "== bench (default)"
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
cat $OUT/bench.json
echo "== bench --impl reference"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"
cat $OUT/bench_ref.json
SMALL="python bench.py --steps 1 --warmup 3 --probes 64 --no-cpu-baseline --no-e2e --no-configs"
echo "== ncu launch list"
$SMALL > $OUT/plain_small.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 172 -c 58 --csv --log-file $OUT/launches.csv $SMALL > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
echo "== ncu --set full (the fused step kernel; k_axpy only exists with SLQ_FUSED=0)"
$SMALL > $OUT/plain_small2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_stepA_tma|k_stepA_csr|k_axpy' -s 160 -c 2 -o $OUT/prof_step $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
