"""Runs BASELINE.json's five configs at FULL size on one B200 through the public API / the device engine, with
size-independent property checks (they are parity-test cases, not bench lines; bench.py times C2).

  python scripts/run_configs.py [--only C1,C3] [--c3-scale 24] [--out gpurun_out/configs.jsonl]

Synthetic inputs are generated on the GPU with torch (input synthesis only; the product path is libslq).
Checks used (no oracle at these sizes):
  * f = identity: the Lanczos quadrature is EXACT for x^T A x (alpha_0 ||x||^2), so quad(identity) must equal an
    independently computed x^T (A x) -- exercises probes, SpMM, recurrence, Golub-Welsch and f at full size;
  * sum(tau) = 1 and theta inside the Gershgorin interval;
  * determinism: the same probe range twice gives bit-identical samples;
  * config-specific known answers (numerical rank of a rank-r Gram matrix, tr(A) by Hutchinson vs the exact trace,
    A (A^-1 x) = x for the f(A) X engine).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _sync_ms(fn):
	import torch

	torch.cuda.synchronize()
	e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
	e0.record()
	out = fn()
	e1.record()
	torch.cuda.synchronize()
	return out, e0.elapsed_time(e1)


def identity_check(op, dtype, k=32, seed=3):
	"""max relative error of quad(identity) against x^T (A x) on one Rademacher block, plus sum(tau) - 1."""
	import torch

	from primate_b200 import engine
	from primate_b200.random import ProbeStream

	n = op.n
	stream = ProbeStream(n, "rademacher", seed=seed, dtype=dtype)
	X = stream.block(k, 0)
	Y = op.matmat(X)
	ref = (X.double() * Y.double()).sum(dim=0)
	out = op.lanczos(stream.block(k, 0), 8, 1e-8, 0, 2)
	theta, tau, y, _ = engine.quadrature(out["alpha"], out["beta"], 8, 0, None, scale=out["xnorm2"])
	rel = ((y - ref).abs() / ref.abs().clamp_min(1e-300)).max().item()
	return {"quad_identity_vs_xtAx_rel": rel, "sum_tau_minus_1": (tau.sum(dim=1) - 1).abs().max().item()}


def slq_run(M, dtype, probes, kb, seed=0):
	"""hutch-style pass over `probes` Rademacher probes: samples (device), ms, probe-steps/s; run twice for determinism."""
	from primate_b200.random import ProbeStream

	stream = ProbeStream(M.shape[0], "rademacher", seed=seed, dtype=dtype)
	M.quad_range(stream, 0, min(probes, kb), kb)  # warm-up (workspace, kernels)
	y1, ms = _sync_ms(lambda: M.quad_range(stream, 0, probes, kb))
	y2, _ = _sync_ms(lambda: M.quad_range(stream, 0, probes, kb))
	return y1, ms, bool((y1 == y2).all().item())


# ---- synthetic operators built on the GPU -------------------------------------------------------------------------
def rmat_laplacian(scale: int, edge_factor: int = 16, seed: int = 0, abcd=(0.57, 0.19, 0.19, 0.05)):
	"""Graph Laplacian L = D - W of a symmetrised, de-duplicated R-MAT graph as device CSR (fp32)."""
	import torch

	a, b, c, _ = abcd
	n, m = 1 << scale, edge_factor << scale
	g = torch.Generator(device="cuda").manual_seed(seed)
	src = torch.zeros(m, dtype=torch.int64, device="cuda")
	dst = torch.zeros(m, dtype=torch.int64, device="cuda")
	for bit in range(scale):
		r = torch.rand(m, generator=g, device="cuda")
		src |= (r >= a + b).long() << bit
		dst |= (((r >= a) & (r < a + b)) | (r >= a + b + c)).long() << bit
		del r
	keep = src != dst
	src, dst = src[keep], dst[keep]
	keys = torch.unique(torch.cat([src * n + dst, dst * n + src]))  # sorted, de-duplicated directed edges
	del src, dst, keep
	deg = torch.bincount(keys // n, minlength=n)
	diag = torch.arange(n, device="cuda", dtype=torch.int64) * (n + 1)
	allk, order = torch.sort(torch.cat([keys, diag]))
	vals = torch.cat([torch.full((keys.numel(),), -1.0, dtype=torch.float32, device="cuda"), deg.float()])[order]
	del keys, diag, order
	rowptr = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
	rowptr[1:] = torch.cumsum(torch.bincount(allk // n, minlength=n), 0)
	colidx = (allk % n).int()
	return rowptr.int(), colidx, vals, n, int(deg.max().item()), float(deg.sum().item())


def laplacian_3d(g: int, dtype):
	"""7-point Dirichlet Laplacian on a g^3 grid as device CSR with sorted columns (C5)."""
	import torch

	n = g**3
	idx = torch.arange(n, device="cuda", dtype=torch.int64)
	x, y, z = idx % g, (idx // g) % g, idx // (g * g)
	offs = torch.tensor([-g * g, -g, -1, 0, 1, g, g * g], device="cuda", dtype=torch.int64)
	valid = torch.stack([z > 0, y > 0, x > 0, torch.ones_like(x, dtype=torch.bool), x < g - 1, y < g - 1, z < g - 1], dim=1)
	cols = (idx[:, None] + offs[None, :])[valid]
	vals = torch.tensor([-1, -1, -1, 6, -1, -1, -1], device="cuda", dtype=dtype).expand(n, 7)[valid]
	rowptr = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
	rowptr[1:] = torch.cumsum(valid.sum(dim=1), 0)
	return rowptr.int(), cols.int(), vals.contiguous(), n


def ritz_max(op, dtype, deg=24):
	"""Largest Ritz value of a short Lanczos run (an estimate of lambda_max from below, within a few %)."""
	from primate_b200 import engine
	from primate_b200.random import ProbeStream

	X = ProbeStream(op.n, "rademacher", seed=99, dtype=dtype).block(2, 0)
	out = op.lanczos(X, deg, 1e-8, 0, 2)
	theta, _, _, _ = engine.quadrature(out["alpha"], out["beta"], deg, -1)
	return float(theta.max().item())


# ---- the five configs -------------------------------------------------------------------------------------------------
def stats(y):
	y = y.double()
	return {"mean": y.mean().item(), "std": y.std().item(), "count": int(y.numel())}


def c1(args):
	"""C1: hutch log-det of a dense 1000x1000 SPD matrix, deg=20, 200 Rademacher probes, fp64 (public API)."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	rng = np.random.default_rng(0)
	B = rng.standard_normal((1000, 1000))
	A = B @ B.T / 1000 + 0.1 * np.eye(1000)
	M = MatrixFunction(A, "log", deg=20, orth=0)
	hutch(M, converge="count", count=200, seed=1)
	t0 = time.perf_counter()
	est = hutch(M, converge="count", count=200, seed=1)
	dt = time.perf_counter() - t0
	exact = float(np.linalg.slogdet(A)[1])
	y, _, _ = slq_run(M, np.float64, 200, 32, seed=1)
	se = y.std().item() / np.sqrt(200)
	assert abs(y.mean().item() - est) < 1e-9 * abs(est), "hutch() and the raw sample mean disagree"
	out = {"config": "C1", "n": 1000, "estimate": float(est), "exact_logdet": exact, "abs_err": abs(est - exact), "std_err": se,
	       "probe_steps_per_s": 200 * 20 / dt, "wall_s": dt}  # fmt: skip
	out.update(identity_check(engine.DeviceOperator(A), np.float64))
	out["ok"] = bool(out["quad_identity_vs_xtAx_rel"] < 1e-10 and out["abs_err"] < max(6 * se, 0.02 * abs(exact)))
	return out


def c2(args):
	"""C2: 2-D Laplacian 1000^2, f=log, deg=50, fp64 (1024 of the 10k probes here; bench.py runs the full job)."""
	import scipy.sparse as sp

	from primate_b200.operators import MatrixFunction

	g = 1000
	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
	A = (sp.kron(sp.identity(g), T) + sp.kron(T, sp.identity(g))).tocsr()
	M = MatrixFunction(A, "log", deg=50, orth=0)
	y, ms, same = slq_run(M, np.float64, 1024, 32)
	out = {"config": "C2", "n": g * g, "nnz": int(A.nnz), "probes": 1024, "deg": 50, "ms": ms, "probe_steps_per_s": 1024 * 50 / ms * 1e3,
	       "deterministic": same, **stats(y)}  # fmt: skip
	out.update(identity_check(M.device_op, np.float64))
	Mi = MatrixFunction(A, "identity", deg=4, orth=0)
	yi, _, _ = slq_run(Mi, np.float64, 256, 32)
	out["trace_identity_z"] = abs(yi.mean().item() - 4.0 * g * g) / (yi.std().item() / 16.0)
	out["ok"] = bool(same and out["quad_identity_vs_xtAx_rel"] < 1e-10 and out["trace_identity_z"] < 5)
	return out


def c3(args):
	"""C3: heat-kernel trace tr(exp(-t L)) of an R-MAT graph Laplacian, scale 24 / 256M edges, fp32, deg=30, 256 probes."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	t0 = time.perf_counter()
	rowptr, colidx, vals, n, dmax, trace = rmat_laplacian(args.c3_scale)
	gen_s = time.perf_counter() - t0
	t0 = time.perf_counter()
	op = engine.DeviceOperator.from_device_csr(rowptr, colidx, vals, n, dtype=np.float32, reorder=args.c3_reorder)
	del rowptr, colidx, vals
	ingest_s = time.perf_counter() - t0
	lam = ritz_max(op, np.float32)
	M = MatrixFunction(op, "exp", deg=30, orth=0, dtype=np.float32, t=-1.0 / lam)
	probes = args.c3_probes
	kb = int(getattr(args, "c3_block", 64))  # fp32 rows of 64 probes (256 B): 32% cheaper per probe than 32 (profiles/r02_c3_experiments.md)
	y, ms, same = slq_run(M, np.float32, probes, kb)
	out = {"config": "C3", "n": n, "nnz": op.nnz, "max_degree": dmax, "lambda_max_ritz": lam, "gen_s": gen_s, "ingest_s": ingest_s, "reorder": str(args.c3_reorder), "probes": probes, "deg": 30,
	       "block": kb, "ms": ms, "probe_steps_per_s": probes * 30 / ms * 1e3, "deterministic": same, **stats(y)}  # fmt: skip
	s = 4
	out["algorithmic_GBps"] = ((op.nnz * (s + 4) + (n + 1) * 4) / kb + 6 * n * s) * probes * 30 / ms / 1e6
	out.update(identity_check(op, np.float32))
	Mi = MatrixFunction(op, "identity", deg=4, orth=0, dtype=np.float32)
	yi, _, _ = slq_run(Mi, np.float32, 64, 32)
	out["trace_identity_z"] = abs(yi.double().mean().item() - trace) / (yi.double().std().item() / 8.0)
	out["ok"] = bool(same and out["quad_identity_vs_xtAx_rel"] < 1e-4 and out["trace_identity_z"] < 5 and 0 < out["mean"] <= n)
	return out


def c4(args):
	"""C4: numerical rank of a dense 32k x 32k Gram matrix (rank 8192), smoothstep, deg=100, full re-orthogonalisation."""
	import torch

	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	n, r = args.c4_n, args.c4_n // 4
	g = torch.Generator(device="cuda").manual_seed(0)
	X = torch.randn((n, r), generator=g, device="cuda", dtype=torch.float64)
	G = (X @ X.T) / r
	G = (G + G.T) / 2
	del X
	op = engine.DeviceOperator(G, dtype=np.float64)
	lam = ritz_max(op, np.float64)
	M = MatrixFunction(op, "smoothstep", deg=100, orth=100, a=1e-6 * lam, b=1e-3 * lam)
	probes = 64
	y, ms, same = slq_run(M, np.float64, probes, args.c4_block)
	st = stats(y)
	out = {"config": "C4", "n": n, "rank": r, "lambda_max_ritz": lam, "probes": probes, "block": args.c4_block, "deg": 100, "orth": 100, "ms": ms,
	       "probe_steps_per_s": probes * 100 / ms * 1e3, "deterministic": same, **st}  # fmt: skip
	out["matrix_GBps"] = n * n * 8 * (probes / args.c4_block) * 100 / ms / 1e6
	out.update(identity_check(op, np.float64))
	out["rank_z"] = abs(st["mean"] - r) / (st["std"] / np.sqrt(probes))
	out["ok"] = bool(same and out["quad_identity_vs_xtAx_rel"] < 1e-10 and out["rank_z"] < 5)
	return out


def c5(args):
	"""C5: 3-D 7-point Laplacian 256^3 (n=16.7M), f=inv, deg=80: SLQ samples and the f(A) X engine at full size; the
	Hutch++ / xdiag drivers (host-side small algebra) on a 96^3 grid through the public API."""
	import torch

	from primate_b200 import engine
	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import ProbeStream
	from primate_b200.trace import hutchpp

	g = args.c5_grid
	t0 = time.perf_counter()
	rowptr, colidx, vals, n = laplacian_3d(g, torch.float64)
	op = engine.DeviceOperator.from_device_csr(rowptr, colidx, vals, n, dtype=np.float64, reorder=True)
	ingest_s = time.perf_counter() - t0
	M = MatrixFunction(op, "inv", deg=80, orth=0)
	probes, kb = 64, 16
	y, ms, same = slq_run(M, np.float64, probes, kb)
	out = {"config": "C5", "n": n, "nnz": op.nnz, "ingest_s": ingest_s, "probes": probes, "deg": 80, "ms": ms,
	       "probe_steps_per_s": probes * 80 / ms * 1e3, "deterministic": same, **stats(y)}  # fmt: skip
	out["algorithmic_GBps"] = ((op.nnz * 12 + (n + 1) * 4) / kb + 6 * n * 8) * probes * 80 / ms / 1e6
	out.update(identity_check(op, np.float64, k=16))
	# f(A) X engine at full size: with f = identity the Krylov combination must reproduce A X.  orth = 0 runs the
	# basis-free two-pass engine (blocks of 32 columns, five panels); the stored-basis formula (blocks of 4: a 43 GB basis)
	# is timed beside it on the same 64 columns.
	from primate_b200 import operators as _ops

	Mid = MatrixFunction(op, "identity", deg=80, orth=0, basis="twopass")
	X = ProbeStream(n, "rademacher", seed=5, dtype=np.float64).block(32, 0)
	ref = op.matmat(X)
	(Y, ms_f) = _sync_ms(lambda: Mid.matmat_block(X))
	out["funm_identity_rel"] = ((Y - ref).abs().max() / ref.abs().max()).item()
	out["funm_block_ms"] = ms_f
	out["funm_block_k"] = 32
	del Y, ref
	X64 = torch.cat([X, ProbeStream(n, "rademacher", seed=5, dtype=np.float64).block(32, 32)], dim=1)
	del X
	_ops._FORCE_TWOPASS = True
	try:
		M.matmat_device(X64[:, :32].contiguous())  # warm-up
		(Y2, out["funm_64col_ms_twopass"]) = _sync_ms(lambda: M.matmat_device(X64))
	finally:
		_ops._FORCE_TWOPASS = False
	_ops._FORCE_STORED_BASIS = True
	try:
		(Y1, out["funm_64col_ms_stored"]) = _sync_ms(lambda: M.matmat_device(X64))
	finally:
		_ops._FORCE_STORED_BASIS = False
	out["funm_twopass_vs_stored_rel"] = ((Y2 - Y1).abs().max() / Y1.abs().max()).item()
	del Y1, Y2, X64
	torch.cuda.empty_cache()
	# drivers on a smaller grid through the public API (host QR / leave-one-out algebra)
	gs = args.c5_small
	rp, ci, va, ns = laplacian_3d(gs, torch.float64)
	ops = engine.DeviceOperator.from_device_csr(rp, ci, va, ns, dtype=np.float64, reorder=True)
	Ms = MatrixFunction(ops, "inv", deg=80, orth=0)
	t0 = time.perf_counter()
	hpp = hutchpp(Ms, m=96, seed=0)
	out["hutchpp_small_s"] = time.perf_counter() - t0
	t0 = time.perf_counter()
	d = xdiag(Ms, m=64, seed=0)
	out["xdiag_small_s"] = time.perf_counter() - t0
	out["hutchpp_small"] = float(hpp)
	out["xdiag_small_sum"] = float(np.sum(d))
	out["hutchpp_vs_xdiag_rel"] = abs(float(hpp) - float(np.sum(d))) / abs(float(hpp))
	out["ok"] = bool(same and out["quad_identity_vs_xtAx_rel"] < 1e-10 and out["funm_identity_rel"] < 1e-9 and
	                 out["funm_twopass_vs_stored_rel"] < 1e-11 and n / 12.0 < out["mean"] and out["hutchpp_vs_xdiag_rel"] < 0.05)  # fmt: skip
	return out


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--only", default="C1,C2,C3,C4,C5")
	ap.add_argument("--c3-scale", type=int, default=24)
	ap.add_argument("--c3-probes", type=int, default=256)
	ap.add_argument("--c3-block", type=int, default=64)
	ap.add_argument("--c3-reorder", default="balance", type=lambda v: False if v in ("0", "none", "False") else v)
	ap.add_argument("--c4-n", type=int, default=32768)
	ap.add_argument("--c4-block", type=int, default=32)
	ap.add_argument("--c5-grid", type=int, default=256)
	ap.add_argument("--c5-small", type=int, default=96)
	ap.add_argument("--out", default="gpurun_out/configs.jsonl")
	args = ap.parse_args()
	import torch

	os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
	fns = {"C1": c1, "C2": c2, "C3": c3, "C4": c4, "C5": c5}
	with open(args.out, "a") as f:
		for name in args.only.split(","):
			t0 = time.perf_counter()
			try:
				res = fns[name](args)
			except Exception as exc:  # noqa: BLE001  (report and go on to the next config)
				import traceback

				traceback.print_exc()
				res = {"config": name, "ok": False, "error": repr(exc)}
			res["total_s"] = time.perf_counter() - t0
			print(json.dumps(res), flush=True)
			f.write(json.dumps(res) + "\n")
			f.flush()
			torch.cuda.empty_cache()


if __name__ == "__main__":
	main()
