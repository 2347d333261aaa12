"""Every libslq kernel family once at small sizes, each result checked against NumPy / a dense eigendecomposition:
CSR step kernels (generic, bulk-copy lean, non-zero-balanced; fp64 / fp32; orth 0 / 5 / deg), dense kernels (padded +
split-K, tensor-core), the two-pass and stored-basis f(A) X engines, the QL warp kernel (eigh / eigvalsh / tqli / fttr),
the probe generators (Rademacher, ziggurat normals, sphere; random access) and the panel kernels with the three block
drivers on top.  36 s on a B200 (profiles/r02c_kernel_families_check.log).

The sizes are chosen so that the script can also run under `compute-sanitizer --tool memcheck|racecheck|synccheck`
(`--skip-tma` leaves out the bulk-copy / mbarrier kernels, whose async-proxy writes racecheck does not model) on a
machine where that tool is available; it is closed on this pool, so here the script is a plain correctness pass."""
import argparse
import os
import sys

import numpy as np
import scipy.sparse as sp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def laplacian_2d(g, dtype=np.float64):
	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g), dtype=dtype)
	I = sp.identity(g, dtype=dtype)  # noqa: E741
	return (sp.kron(I, T) + sp.kron(T, I)).tocsr().astype(dtype)


def hub_matrix(n, hub_nnz, seed=0):
	"""Symmetric CSR with one row of `hub_nnz` non-zeros (the non-zero-balanced step kernel) on a diagonally dominant band."""
	rng = np.random.default_rng(seed)
	A = sp.diags([1.0, 4.0, 1.0], [-1, 0, 1], shape=(n, n)).tolil()
	cols = rng.choice(np.arange(1, n), size=hub_nnz, replace=False)
	for c in cols:
		A[0, c] = A[c, 0] = 0.01
	A[0, 0] = 4.0 + 0.01 * hub_nnz
	return A.tocsr()


def exact_quad(A, x, f):
	ew, ev = np.linalg.eigh(A.toarray() if sp.issparse(A) else A)
	c = ev.T @ x
	return float(np.sum(f(ew) * c * c))


def case_csr(tma: bool):
	import torch

	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	for dtype, tol in ((np.float64, 1e-8), (np.float32, 2e-3)):
		A = laplacian_2d(24, dtype) + sp.identity(576, dtype=dtype)
		n = A.shape[0]
		rng = np.random.default_rng(1)
		for k in ((1, 4, 8, 64) if tma else (1, 4)):  # k >= 8 takes the TMA step kernel, below it the generic CSR one
			X = rng.standard_normal((n, k)).astype(dtype)
			for orth in (0, 5, 40):
				M = MatrixFunction(A, "log", deg=40, orth=orth, dtype=dtype)
				y = M.quad_device(torch.from_numpy(X).cuda()).double().cpu().numpy()
				ref = np.array([exact_quad(A.astype(np.float64), X[:, j].astype(np.float64), np.log) for j in range(k)])
				assert np.max(np.abs(y / ref - 1)) < tol, ("csr", dtype, k, orth, np.max(np.abs(y / ref - 1)))
	M = MatrixFunction(laplacian_2d(24) + sp.identity(576), "log", deg=20, orth=0)
	est = hutch(M, converge="count", count=40, seed=3)
	assert np.isfinite(est)
	if tma:
		H = hub_matrix(3000, 700)
		X = np.random.default_rng(2).standard_normal((3000, 16))
		M = MatrixFunction(H, "sqrt", deg=30, orth=0)
		y = M.quad_device(torch.from_numpy(X).cuda()).cpu().numpy()
		ref = np.array([exact_quad(H, X[:, j], np.sqrt) for j in range(16)])
		assert np.max(np.abs(y / ref - 1)) < 1e-6, ("hub", np.max(np.abs(y / ref - 1)))
	print("csr ok", flush=True)


def case_dense(tma: bool):
	import torch

	from primate_b200.operators import MatrixFunction

	rng = np.random.default_rng(4)
	for n in ((150, 1000) if tma else (150,)):
		B = rng.standard_normal((n, n))
		A = B @ B.T / n + 0.5 * np.eye(n)
		for dtype, tol in ((np.float64, 1e-8), (np.float32, 2e-3)):
			for k in ((1, 8, 32) if tma else (1,)):
				X = rng.standard_normal((n, k)).astype(dtype)
				M = MatrixFunction(A.astype(dtype), "log", deg=min(n, 30), orth=3, dtype=dtype)
				y = M.quad_device(torch.from_numpy(X).cuda()).double().cpu().numpy()
				ref = np.array([exact_quad(A, X[:, j].astype(np.float64), np.log) for j in range(k)])
				assert np.max(np.abs(y / ref - 1)) < max(tol, 1e-4), ("dense", n, dtype, k, np.max(np.abs(y / ref - 1)))
	print("dense ok", flush=True)


def case_funm(tma: bool):
	import torch

	from primate_b200.operators import MatrixFunction

	A = laplacian_2d(20) + sp.identity(400)
	ew, ev = np.linalg.eigh(A.toarray())
	fA = (ev * np.exp(-0.3 * ew)) @ ev.T
	rng = np.random.default_rng(5)
	for k in ((1, 4, 8, 37) if tma else (1, 4)):
		X = rng.standard_normal((400, k))
		for basis in ("twopass", "stored"):
			M = MatrixFunction(A, "exp", deg=40, orth=0, t=-0.3, basis=basis)
			Y = M.matmat_device(torch.from_numpy(X).cuda()).cpu().numpy()
			assert np.max(np.abs(Y - fA @ X)) < 1e-8 * np.max(np.abs(Y)), ("funm", k, basis)
	print("funm ok", flush=True)


def case_tridiag():
	from primate_b200.fttr import fttr
	from primate_b200.tridiag import eigh_tridiag, eigvalsh_tridiag, tqli

	rng = np.random.default_rng(6)
	for m in (1, 2, 20, 33, 150):
		d, e = rng.uniform(size=m), rng.uniform(size=max(m - 1, 0))
		T = np.diag(d) + np.diag(e, 1) + np.diag(e, -1)
		ew = np.linalg.eigvalsh(T)
		w, V = eigh_tridiag(d, e)
		assert np.max(np.abs(np.sort(w) - ew)) < 1e-12 and np.max(np.abs(V.T @ V - np.eye(m))) < 1e-12, m
		assert np.max(np.abs(np.sort(eigvalsh_tridiag(d, e)) - ew)) < 1e-12
		dd, ee, Z = d.copy(), np.append(0.0, e), np.eye(m)
		tqli(dd, ee, Z, 60)
		assert np.max(np.abs(np.sort(dd) - ew)) < 1e-12, m
		if m >= 2:
			wts = np.zeros(m)
			fttr(np.sort(w), d, np.append(0.0, e), m, wts)
			assert np.all(np.isfinite(wts)) and np.all(wts > 0)
	print("tridiag ok", flush=True)


def case_probes():
	from primate_b200.random import ProbeStream

	n = 20011
	for pdf in ("rademacher", "normal", "sphere"):
		S = ProbeStream(n, pdf, seed=11, dtype=np.float64)
		X = S.block(5, 0).cpu().numpy()
		X2 = S.block(3, 5).cpu().numpy()
		R = ProbeStream(n, pdf, seed=11).block(3, 5).cpu().numpy()  # random access: a scan from position 0
		assert np.array_equal(X2, R)
		rng = np.random.default_rng(11)
		W = np.empty((n, 8), order="F")
		if pdf == "rademacher":
			for j in range(8):
				W[:, j] = np.where(rng.random(n) >= 0.5, 1.0, -1.0)
		else:
			for j in range(8):
				W[:, j] = rng.standard_normal(n)
				if pdf == "sphere":
					W[:, j] *= np.sqrt(n) / np.sqrt(np.sum(W[:, j : j + 1] ** 2, axis=0))
		got = np.concatenate([X, X2], axis=1)
		if pdf == "sphere":
			assert np.max(np.abs(got - W)) < 1e-13, pdf  # (the exact NumPy op order is pinned by tests/test_gpu_normals.py)
		else:
			assert np.array_equal(got, W), pdf
	print("probes ok", flush=True)


def case_panels_and_drivers():
	import torch

	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction
	from primate_b200.panels import DevicePanels
	from primate_b200.random import ProbeStream
	from primate_b200.trace import hutchpp, xtrace

	n = 1203
	A = sp.diags(np.linspace(1.0, 2.0, n)).tocsr()
	rng = np.random.default_rng(7)
	for dtype in (np.float64, np.float32):
		M = MatrixFunction(A.astype(dtype), "identity", deg=4, orth=0, dtype=dtype)
		P = DevicePanels(M, ProbeStream(n, "rademacher", seed=0, dtype=dtype))
		for a, b in ((7, 7), (32, 32), (96, 104), (130, 40)):
			X, Y = rng.standard_normal((n, a)).astype(dtype), rng.standard_normal((n, b)).astype(dtype)
			Xd, Yd = torch.from_numpy(X).cuda(), torch.from_numpy(Y).cuda()
			tol = 1e-12 if dtype == np.float64 else 1e-5
			assert np.max(np.abs(P.gram(Xd, Yd) - X.astype(np.float64).T @ Y.astype(np.float64))) < tol * n
			S = rng.standard_normal((a, b))
			assert np.max(np.abs(P.matmul(Xd, S).double().cpu().numpy() - X.astype(np.float64) @ S)) < tol * 100 * a
			if a == b:
				assert np.max(np.abs(P.coldot(Xd, Yd) - np.einsum("ij,ij->j", X.astype(np.float64), Y.astype(np.float64)))) < tol * n
				P.rowdot(Xd, Yd)
		Q, R = P.qr(torch.from_numpy(rng.standard_normal((n, 24)).astype(dtype)).cuda())
		G = P.gram(Q, Q)
		assert np.max(np.abs(G - np.eye(24))) < (1e-13 if dtype == np.float64 else 1e-5)
	L = laplacian_2d(16) + sp.identity(256)
	M = MatrixFunction(L, "exp", deg=30, orth=0, t=-0.2)
	tr = float(np.sum(np.exp(-0.2 * np.linalg.eigvalsh(L.toarray()))))
	assert abs(hutchpp(M, m=96, seed=1) / tr - 1) < 5e-2
	assert abs(xtrace(M, batch=32, seed=2) / tr - 1) < 5e-2
	d = xdiag(M, 60, seed=3)
	assert np.all(np.isfinite(np.asarray(d)))
	print("panels + drivers ok", flush=True)


if __name__ == "__main__":
	ap = argparse.ArgumentParser()
	ap.add_argument("--skip-tma", action="store_true", help="leave out the bulk-copy / mbarrier kernels (racecheck)")
	ap.add_argument("--only", default="")
	a = ap.parse_args()
	tma = not a.skip_tma
	cases = {
		"csr": lambda: case_csr(tma),
		"dense": lambda: case_dense(tma),
		"funm": lambda: case_funm(tma),
		"tridiag": case_tridiag,
		"probes": case_probes,
		"panels": case_panels_and_drivers,
	}
	for name, fn in cases.items():
		if not a.only or name in a.only.split(","):
			fn()
	print("ALL CASES OK")
