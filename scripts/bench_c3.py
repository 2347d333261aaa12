"""Experiments on the irregular SpMM (config C3: R-MAT graph Laplacian, fp32): times slq_op_matmat (the plain SpMM of the
step-A kernel) with CUDA events under different row orders / block widths, and on a copy of the matrix whose column
indices are folded into a small range (every gather an L2 hit) to separate memory-system time from kernel overhead.
  python scripts/bench_c3.py [--scale 22]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--scale", type=int, default=22)
	ap.add_argument("--ks", default="32")
	ap.add_argument("--fold", type=int, default=1, help="also time the folded-column copy")
	args = ap.parse_args()
	import torch

	from primate_b200 import engine
	from scripts.run_configs import rmat_laplacian

	rowptr, colidx, vals, n, dmax, _ = rmat_laplacian(args.scale)
	nnz = int(colidx.numel())

	def timed(op, k, reps=5):
		X = torch.randn((n, k), device="cuda", dtype=torch.float32)
		for _ in range(2):
			op.matmat(X)
		torch.cuda.synchronize()
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for _ in range(reps):
			op.matmat(X)
		e1.record()
		torch.cuda.synchronize()
		return e0.elapsed_time(e1) / reps

	for k in [int(x) for x in args.ks.split(",")]:
		for order in (False, "balance"):
			op = engine.DeviceOperator.from_device_csr(rowptr, colidx, vals, n, dtype=np.float32, reorder=order)
			ms = timed(op, k)
			alg = (nnz * 8 + (n + 1) * 4 + 2 * n * k * 4) / ms / 1e6
			print(json.dumps({"scale": args.scale, "k": k, "order": str(order), "case": "graph", "ms": ms, "nnz_per_ns": nnz / ms / 1e6,
			                  "gather_GBps": nnz * k * 4 / ms / 1e6, "algorithmic_GBps": alg}), flush=True)  # fmt: skip
			del op
		if args.fold:
			for bits in (12, 18):
				folded = (colidx & ((1 << bits) - 1)).contiguous()
				for order in (False, "balance"):  # ("balance" renumbers the folded columns again, but they stay inside a 2^bits-row set)
					op = engine.DeviceOperator.from_device_csr(rowptr, folded, vals, n, dtype=np.float32, reorder=order)
					ms = timed(op, k)
					print(json.dumps({"scale": args.scale, "k": k, "order": str(order), "case": f"columns folded to {1 << bits} rows ({(1 << bits) * k * 4 >> 10} KB)",
					                  "ms": ms, "nnz_per_ns": nnz / ms / 1e6, "gather_GBps": nnz * k * 4 / ms / 1e6}), flush=True)  # fmt: skip
					del op
				del folded


if __name__ == "__main__":
	main()
