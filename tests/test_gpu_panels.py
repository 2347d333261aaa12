"""The tall-skinny panel kernels behind the block drivers (libslq: slq_panel_gram / _matmul / _rowdot / _coldot) against
NumPy, CholeskyQR2 built on them against np.linalg.qr, and XTrace with device panels against the reference's host
algebra (src/primate/trace.py:185-315) on the same probes."""
import numpy as np
import pytest
import scipy.sparse as sp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
	"""DevicePanels on a small MatrixFunction (the operator is only needed for shapes / dtype here)."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.panels import DevicePanels
	from primate_b200.random import ProbeStream

	def make(n, dtype=np.float64):
		A = sp.diags(np.linspace(1.0, 2.0, n)).tocsr().astype(dtype)
		M = MatrixFunction(A, "identity", deg=4, orth=0, dtype=dtype)
		return DevicePanels(M, ProbeStream(n, "rademacher", seed=0, dtype=dtype))

	return make


@pytest.mark.parametrize("n,a,b", [(1000, 7, 7), (5003, 37, 96), (2048, 130, 200), (333, 1, 64), (40000, 96, 96), (3001, 32, 32), (7001, 64, 200),
                                   (515, 96, 104), (17, 4, 8)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_gram_matmul_dots_vs_numpy(P, n, a, b, dtype):
	import torch

	rng = np.random.default_rng(n + a)
	X, Y = rng.standard_normal((n, a)).astype(dtype), rng.standard_normal((n, b)).astype(dtype)
	Xd, Yd = torch.from_numpy(X).cuda(), torch.from_numpy(Y).cuda()
	p = P(n, dtype)
	tol = 1e-12 if dtype == np.float64 else 1e-5
	G = p.gram(Xd, Yd)
	ref = X.astype(np.float64).T @ Y.astype(np.float64)
	assert G.shape == (a, b) and np.max(np.abs(G - ref)) < tol * n
	assert np.array_equal(G, p.gram(Xd, Yd))  # deterministic
	S = rng.standard_normal((a, b))
	Z = p.matmul(Xd, S).cpu().numpy()
	assert np.max(np.abs(Z - X.astype(np.float64) @ S)) < (1e-12 if dtype == np.float64 else 1e-4) * a
	Y2 = Yd.clone()
	p.axpy_matmul(Y2, Xd, S)
	assert np.max(np.abs(Y2.cpu().numpy() - (Y.astype(np.float64) - X.astype(np.float64) @ S))) < (1e-12 if dtype == np.float64 else 1e-4) * a
	if a == b:
		rd = p.rowdot(Xd, Yd).cpu().numpy()
		assert np.max(np.abs(rd - np.einsum("ij,ij->i", X.astype(np.float64), Y.astype(np.float64)))) < tol * a
		cd = p.coldot(Xd, Yd)
		assert np.max(np.abs(cd - np.einsum("ij,ij->j", X.astype(np.float64), Y.astype(np.float64)))) < tol * n


def test_panels_with_row_stride_views(P):
	"""Column ranges of a wider buffer are panels (row stride = buffer width), as XTrace uses them."""
	import torch

	n = 3001
	rng = np.random.default_rng(1)
	W = rng.standard_normal((n, 50))
	Wd = torch.from_numpy(W).cuda()
	p = P(n)
	A, B = Wd[:, 3:20], Wd[:, 20:45]
	assert np.max(np.abs(p.gram(A, B) - W[:, 3:20].T @ W[:, 20:45])) < 1e-10
	S = rng.standard_normal((17, 5))
	out = torch.zeros((n, 50), dtype=torch.float64, device="cuda")
	p._matmul(A, S, 1.0, 0.0, out[:, 10:15])
	assert np.max(np.abs(out.cpu().numpy()[:, 10:15] - W[:, 3:20] @ S)) < 1e-11
	assert float(out[:, :10].abs().max()) == 0.0 and float(out[:, 15:].abs().max()) == 0.0
	assert np.max(np.abs(p.rowdot(A, Wd[:, 30:47]).cpu().numpy() - np.einsum("ij,ij->i", W[:, 3:20], W[:, 30:47]))) < 1e-11


@pytest.mark.parametrize("m,cond", [(8, 1e2), (64, 1e4), (96, 1e6)])
def test_cholqr2_orthonormal_and_reconstructs(P, m, cond):
	import torch

	n = 20000
	rng = np.random.default_rng(m)
	U, _ = np.linalg.qr(rng.standard_normal((n, m)))
	V, _ = np.linalg.qr(rng.standard_normal((m, m)))
	Y = (U * np.logspace(0, -np.log10(cond), m)) @ V.T
	p = P(n)
	Q, R = p.qr(torch.from_numpy(Y).cuda())
	Qh = Q.cpu().numpy()
	assert np.allclose(np.triu(R), R) and np.all(np.diag(R) > 0)
	assert np.max(np.abs(Qh.T @ Qh - np.eye(m))) < 1e-13 * m
	assert np.max(np.abs(Qh @ R - Y)) < 1e-13 * np.abs(Y).max() * m
	Qr, Rr = np.linalg.qr(Y)
	sgn = np.sign(np.diag(Rr))
	assert np.max(np.abs(Qh - Qr * sgn)) < 1e-9 * cond / 1e2  # same factorisation up to the column signs


def test_xtrace_device_panels_match_host_algebra(golden):
	"""XTrace on a MatrixFunction: W, Q, A Q on the GPU, sphere probes from the device ziggurat, block CGS2 + CholeskyQR2;
	against the same driver on the operator wrapped as a plain LinearOperator (NumPy panels, np.linalg.qr) -- and both
	equal the reference's golden value when f is the identity at full degree."""
	from scipy.sparse.linalg import LinearOperator

	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import xtrace

	A = golden["A150"]
	M = MatrixFunction(A, "exp", deg=40, orth=40, t=0.5)
	plain = LinearOperator(M.shape, matvec=M._matvec, matmat=M._matmat, dtype=M.dtype)
	for batch in (10, 32, 150):
		g1, g2 = np.random.default_rng(8), np.random.default_rng(8)
		dev, res = xtrace(M, batch=batch, seed=g1, full=True)
		host = xtrace(plain, batch=batch, seed=g2)
		assert abs(dev / host - 1) < 1e-9 and len(res.estimator) == 150
		assert g1.bit_generator.state == g2.bit_generator.state  # the device sphere stream consumed the same draws
	ew = np.linalg.eigvalsh(A)
	assert abs(dev - np.exp(0.5 * ew).sum()) < 1e-3 * np.exp(0.5 * ew).sum()  # tests/test_trace.py:60-89: within 1e-2 of the trace
	Mi = MatrixFunction(A, "identity", deg=150, orth=150)
	assert abs(xtrace(Mi, batch=10, seed=8) - golden["xtrace_A_b10_seed8"].item()) < 1e-8
