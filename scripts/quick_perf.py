"""Scratch perf probe for the Lanczos block kernels on config C2 (not the bench contract; see bench.py)."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--grid", type=int, default=1000)
	ap.add_argument("--k", type=int, default=32)
	ap.add_argument("--deg", type=int, default=50)
	ap.add_argument("--blocks", type=int, default=4)
	ap.add_argument("--orth", type=int, default=0)
	ap.add_argument("--dtype", default="float64")
	args = ap.parse_args()
	import scipy.sparse as sp
	import torch

	from primate_b200 import engine

	dt = np.dtype(args.dtype)
	g = args.grid
	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
	A = (sp.kron(sp.identity(g), T) + sp.kron(T, sp.identity(g))).tocsr().astype(dt)
	n = A.shape[0]
	op = engine.DeviceOperator(A, dtype=dt)
	st, inc = 12345678901234567890123, (987654321 << 1) | 1
	ncv = max(2, args.orth)
	s = dt.itemsize
	bytes_step = (A.nnz * (s + 4) + (n + 1) * 4) / args.k + 6 * n * s  # SURVEY 8(d), per probe-step
	for rep in range(3):
		torch.cuda.synchronize()
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for b in range(args.blocks):
			X = engine.rademacher_block(st, inc, b * args.k, args.k, n, dt)
			out = op.lanczos(X, args.deg, 1e-8, args.orth, ncv)
			theta, tau, y, info = engine.quadrature(out["alpha"], out["beta"], args.deg, 3, None, scale=out["xnorm2"], want_nodes=False)
		e1.record()
		torch.cuda.synchronize()
		ms = e0.elapsed_time(e1)
		ps = args.blocks * args.k * args.deg
		print(f"rep {rep}: {ms:.2f} ms for {ps} probe-steps -> {ps / ms * 1e3:.4g} probe-steps/s; "
		      f"algorithmic {bytes_step * ps / ms / 1e6:.1f} GB/s; y[0]={y[0].item():.6f}", flush=True)  # fmt: skip
	# SpMM alone
	X = engine.rademacher_block(st, inc, 0, args.k, n, dt)
	perm, op.perm = op.perm, None  # time the SpMM kernel alone (in the operator's own row order)
	for rep in range(3):
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for _ in range(20):
			Y = op.matmat(X)
		e1.record()
		torch.cuda.synchronize()
		ms = e0.elapsed_time(e1) / 20
		b = A.nnz * (s + 4) + (n + 1) * 4 + 2 * n * args.k * s
		print(f"spmm: {ms * 1e3:.1f} us, {b / ms / 1e6:.1f} GB/s algorithmic", flush=True)


if __name__ == "__main__":
	main()
