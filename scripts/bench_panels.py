"""Times the panel kernels at config C5's shapes (n = 16.7 M rows, m = 32 / 64 / 96 columns, fp64) with CUDA events.
  python scripts/bench_panels.py [--n 16777216]      SLQ_PANEL=fma selects the scalar panel product (A/B)."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--n", type=int, default=1 << 24)
	args = ap.parse_args()
	import scipy.sparse as sp
	import torch
	from primate_b200.operators import MatrixFunction
	from primate_b200.panels import DevicePanels
	from primate_b200.random import ProbeStream

	n = args.n
	M = MatrixFunction(sp.identity(n, format="csr"), "identity", deg=2, orth=0)
	P = DevicePanels(M, ProbeStream(n, "rademacher", seed=0))

	def timed(fn, reps=3):
		fn()
		torch.cuda.synchronize()
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for _ in range(reps):
			fn()
		e1.record()
		torch.cuda.synchronize()
		return e0.elapsed_time(e1) / reps

	for m in (32, 64, 96):
		X = torch.randn((n, m), device="cuda", dtype=torch.float64)
		Y = torch.randn((n, m), device="cuda", dtype=torch.float64)
		S = np.random.default_rng(0).standard_normal((m, m))
		rec = {"n": n, "m": m, "kernel": os.environ.get("SLQ_PANEL", "dmma"), "panel_GB": n * m * 8 / 1e9}
		rec["matmul_ms"] = timed(lambda: P.matmul(X, S))
		rec["matmul_GBps"] = 2 * n * m * 8 / rec["matmul_ms"] / 1e6
		rec["matmul_tflops"] = 2.0 * n * m * m / rec["matmul_ms"] / 1e9
		rec["gram_ms"] = timed(lambda: P.gram(X, Y))
		rec["gram_tflops"] = 2.0 * n * m * m / rec["gram_ms"] / 1e9
		rec["rowdot_ms"] = timed(lambda: P.rowdot(X, Y))
		rec["coldot_ms"] = timed(lambda: P.coldot(X, Y))
		rec["qr_ms"] = timed(lambda: P.qr(X), reps=1)
		print(json.dumps(rec), flush=True)
		del X, Y


if __name__ == "__main__":
	main()
