"""Times the dense symmetric block product Y = A X (slq_op_matmat, config C4's shape: n = 32768, k = 32, fp64) with CUDA
events, next to cuBLAS DGEMM on the same shape (torch.matmul: a library yardstick, not the product path).
  python scripts/bench_dense.py [--n 32768] [--k 32]        SLQ_DENSE=fma selects the scalar-FMA kernel (A/B)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--n", type=int, default=32768)
	ap.add_argument("--k", type=int, default=32)
	ap.add_argument("--reps", type=int, default=10)
	ap.add_argument("--row-major", action="store_true", help="no band-major copy (DeviceOperator(tiled=False))")
	ap.add_argument("--ld-pad", type=int, default=0, help="row stride of A = n + this many elements (power-of-two strides vs HBM channel hashing)")
	args = ap.parse_args()
	import torch

	from primate_b200 import engine

	n, k = args.n, args.k
	g = torch.Generator(device="cuda").manual_seed(0)
	A = torch.randn((n, n), generator=g, device="cuda", dtype=torch.float64)
	A = (A + A.T) / 2
	if args.ld_pad:
		W = torch.zeros((n, n + args.ld_pad), device="cuda", dtype=torch.float64)
		W[:, :n] = A
		del A
		A = W[:, :n]
	X = torch.randn((n, k), generator=g, device="cuda", dtype=torch.float64)
	op = engine.DeviceOperator(A, dtype=np.float64, tiled=not args.row_major)

	def timed(fn):
		for _ in range(3):
			fn()
		torch.cuda.synchronize()
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for _ in range(args.reps):
			out = fn()
		e1.record()
		torch.cuda.synchronize()
		return out, e0.elapsed_time(e1) / args.reps

	Y, ms = timed(lambda: op.matmat(X))
	Yref, ms_blas = timed(lambda: torch.matmul(A, X))
	err = ((Y - Yref).abs().max() / Yref.abs().max()).item()
	flops = 2.0 * n * n * k
	print(json.dumps({"n": n, "k": k, "ld_pad": args.ld_pad, "tiled": op.tiled, "kernel": os.environ.get("SLQ_DENSE", "dmma"), "ms": ms, "tflops": flops / ms / 1e9,
	                  "cublas_ms": ms_blas, "cublas_tflops": flops / ms_blas / 1e9, "vs_cublas": ms_blas / ms, "rel_err_vs_cublas": err}))  # fmt: skip


if __name__ == "__main__":
	main()
