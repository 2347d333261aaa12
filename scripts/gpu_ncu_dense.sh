#!/bin/bash
# ncu of k_dense_tma on config C4 (fp64, k = 32: FMA-pipe-bound) and torch's (cuBLAS) DGEMM on the same shape as a yardstick.
OUT=gpurun_out/${1:-ncudense}
mkdir -p $OUT
python - > $OUT/dgemm_yardstick.txt 2>&1 <<'PY'
import torch
n, k = 32768, 32
A = torch.randn(n, n, dtype=torch.float64, device="cuda"); X = torch.randn(n, k, dtype=torch.float64, device="cuda")
for _ in range(3): Y = A @ X
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(10): Y = A @ X
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"torch (cuBLAS) fp64 {n}x{n} @ {n}x{k}: {ms:.3f} ms, {2*n*n*k/ms/1e9:.1f} TFLOP/s, {n*n*8/ms/1e6:.0f} GB/s of A")
PY
cat $OUT/dgemm_yardstick.txt
CMD="python scripts/run_configs.py --only C4 --out $OUT/c4.jsonl"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_dense_tma -s 20 -c 2 -o $OUT/prof_dense $CMD > $OUT/ncu.log 2>&1
echo "ncu exit $?"
ncu -i $OUT/prof_dense.ncu-rep --page details > $OUT/ncu_dense_details.txt 2>&1
