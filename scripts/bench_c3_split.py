"""Feasibility experiment for a hubs-first split of irregular operators (config C3): times the existing kernels on the
SHORT-row part (rows of <= 64 non-zeros: the lean step-A variant applies) and on the LONG-row part of an R-MAT
Laplacian separately.   python scripts/bench_c3_split.py [--scale 22] [--k 32]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--scale", type=int, default=22)
	ap.add_argument("--k", type=int, default=32)
	ap.add_argument("--thr", type=int, default=64)
	args = ap.parse_args()
	import torch

	from primate_b200 import engine
	from scripts.run_configs import rmat_laplacian

	rowptr, colidx, vals, n, dmax, _ = rmat_laplacian(args.scale)
	rp = rowptr.long()
	lens = rp[1:] - rp[:-1]
	nnz = int(colidx.numel())
	is_long = lens > args.thr
	row_of = torch.repeat_interleave(torch.arange(n, device="cuda"), lens)
	keep_long = is_long[row_of]

	def sub(mask_nnz):
		ln = torch.zeros(n, dtype=torch.int64, device="cuda").index_add_(0, row_of[mask_nnz], torch.ones(int(mask_nnz.sum()), dtype=torch.int64, device="cuda"))
		p = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
		p[1:] = torch.cumsum(ln, 0)
		return p.int(), colidx[mask_nnz].contiguous(), vals[mask_nnz].contiguous()

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

	out = {"scale": args.scale, "k": args.k, "n": n, "nnz": nnz, "long_rows": int(is_long.sum()), "long_nnz": int(keep_long.sum()),
	       "max_degree": dmax}  # fmt: skip
	full = engine.DeviceOperator.from_device_csr(rowptr, colidx, vals, n, dtype=np.float32, reorder="balance")
	out["full_balance_ms"] = timed(full, args.k)
	del full
	rs, cs, vs = sub(~keep_long)
	short = engine.DeviceOperator.from_device_csr(rs, cs, vs, n, dtype=np.float32, reorder=False)
	out["short_part_lean_ms"] = timed(short, args.k)
	out["short_nnz_per_ns"] = (nnz - out["long_nnz"]) / out["short_part_lean_ms"] / 1e6
	del short
	rl, cl, vl = sub(keep_long)
	for order in (False, "balance"):
		long_ = engine.DeviceOperator.from_device_csr(rl, cl, vl, n, dtype=np.float32, reorder=order)
		out[f"long_part_balanced_kernel_{order}_ms"] = timed(long_, args.k)
		del long_
	print(json.dumps(out), flush=True)


if __name__ == "__main__":
	main()
