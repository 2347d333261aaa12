"""The basis-free two-pass f(A) X engine (libslq: slq_funm_matmat / slq_lanczos_replay) against the stored-basis
formula of the reference (src/primate/operators.py:102-124: y = ||x|| Q Y f(Theta) Y^T e_1) -- same start blocks, same
operator, both through the C ABI -- and against the goldens produced by the unmodified reference."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import laplacian_2d

pytestmark = pytest.mark.gpu


def _both(A, fun, deg, X, dtype=np.float64, **kw):
	"""(two-pass, stored-basis) results of M @ X on the device for the same host panel X."""
	import torch

	from primate_b200.operators import MatrixFunction

	Xd = torch.from_numpy(np.ascontiguousarray(X, dtype=dtype)).cuda()
	y2 = MatrixFunction(A, fun, deg=deg, orth=0, dtype=dtype, basis="twopass", **kw).matmat_device(Xd).cpu().numpy()
	assert np.array_equal(Xd.cpu().numpy(), np.asarray(X, dtype=dtype))  # start panel left intact
	y1 = MatrixFunction(A, fun, deg=deg, orth=0, dtype=dtype, basis="stored", **kw).matmat_device(Xd).cpu().numpy()
	return y2, y1


def _L30(golden):
	return sp.csr_matrix((golden["L30_data"], golden["L30_indices"], golden["L30_indptr"]), shape=(900, 900))


@pytest.mark.parametrize("fun,kw", [("log", {}), ("exp", {"t": -0.5}), ("identity", {})])
def test_twopass_matches_reference_goldens_dense(golden, fun, kw):
	"""orth = 0 goldens of the reference's `_matvec` (dense operator: generic replay path, k = 1)."""
	from primate_b200.operators import MatrixFunction

	A, v = golden["A150"], golden["v42"]
	M = MatrixFunction(A, fun, deg=20, orth=0, basis="twopass", **kw)
	y = np.ravel(M._matvec(v.copy()))
	g = np.ravel(golden[f"matvec_{fun}_orth0_v42"])
	assert np.max(np.abs(y - g)) < 1e-8 * np.max(np.abs(g))
	y2, y1 = _both(A, fun, 20, v[:, None], **kw)
	assert np.max(np.abs(y2 - y1)) <= 1e-12 * np.max(np.abs(y1))


@pytest.mark.parametrize("k", [1, 2, 4, 8, 16, 32, 64, 37])
def test_twopass_equals_stored_basis_csr(golden, k):
	"""CSR operator: k < 8 takes the generic replay (SpMM + replay axpy), k >= 8 the fused one-kernel-per-step replay;
	37 columns = 32 + 4 + 1 blocks.  Equality with the stored-basis path to 1e-12 of the largest entry."""
	A = _L30(golden)
	X = np.random.default_rng(k).standard_normal((A.shape[0], k))
	for fun, kw, deg in (("exp", {"t": -0.3}, 30), ("inv", {}, 50)):
		y2, y1 = _both(A + (0.5 * sp.identity(900) if fun == "inv" else 0 * sp.identity(900)), fun, deg, X, **kw)
		assert np.max(np.abs(y2 - y1)) <= 1e-12 * np.max(np.abs(y1)), (fun, k)


@pytest.mark.parametrize("reorder", [False, True, "balance"])
def test_twopass_long_rows_and_row_orders(reorder):
	"""Hub rows (the non-zero-balanced fused variant) and both internal row orders: results in the caller's order."""
	import torch

	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	rng = np.random.default_rng(2)
	n = 5003
	S = sp.random(n, n, density=4.0 / n, random_state=3, format="lil")
	for hub in (0, 2500, n - 1):
		cols = rng.choice(n, size=900, replace=False)
		S[hub, cols] = rng.standard_normal(900)
	S = sp.csr_matrix(S)
	A = (S + S.T + sp.identity(n) * 50.0).tocsr()
	A.sort_indices()
	assert np.diff(A.indptr).max() > 64
	op = engine.DeviceOperator(A, reorder=reorder)
	X = rng.standard_normal((n, 16))
	Xd = torch.from_numpy(X).cuda()
	y2 = MatrixFunction(op, "sqrt", deg=25, orth=0, basis="twopass").matmat_device(Xd).cpu().numpy()
	y1 = MatrixFunction(op, "sqrt", deg=25, orth=0, basis="stored").matmat_device(Xd).cpu().numpy()
	assert np.max(np.abs(y2 - y1)) <= 1e-12 * np.max(np.abs(y1))
	ew, U = np.linalg.eigh(A.toarray())
	ref = (U * np.sqrt(ew)) @ (U.T @ X[:, :2])
	assert np.max(np.abs(y2[:, :2] - ref)) < 1e-6 * np.max(np.abs(ref))


def test_twopass_early_break_and_host_callable(golden):
	"""A start vector in a 3-dimensional invariant subspace breaks down at step 3 (lanczos.h:140): the replay zeroes the
	steps that never ran, exactly as the reference's zero-filled basis columns do; a Python callable takes the
	host-coefficient replay entry (slq_lanczos_replay)."""
	import torch

	from primate_b200.operators import MatrixFunction

	n = 640
	d = np.repeat([1.0, 2.0, 5.0], [200, 200, 240])
	A = sp.diags(d).tocsr()
	rng = np.random.default_rng(0)
	X = rng.standard_normal((n, 8))
	for fun in ("exp", lambda x: np.exp(x)):
		M = MatrixFunction(A, fun, deg=12, orth=0, basis="twopass")
		Y = M.matmat_device(torch.from_numpy(X).cuda()).cpu().numpy()
		assert np.all(np.isfinite(Y))
		assert np.max(np.abs(Y - np.exp(d)[:, None] * X)) < 1e-9 * np.max(np.abs(Y))
	A30 = _L30(golden)
	V = rng.standard_normal((900, 5))
	y_builtin = MatrixFunction(A30, "exp", deg=30, orth=0, t=-0.3, basis="twopass").matmat_device(torch.from_numpy(V).cuda()).cpu().numpy()
	y_callable = MatrixFunction(A30, lambda x: np.exp(-0.3 * x), deg=30, orth=0, basis="twopass").matmat_device(torch.from_numpy(V).cuda()).cpu().numpy()
	assert np.max(np.abs(y_builtin - y_callable)) <= 1e-11 * np.max(np.abs(y_builtin))


def test_twopass_fp32_and_dense_tiled():
	"""fp32 CSR (fused replay) and a dense operator with n % 16 == 0, k = 32 (tiled dense product + replay axpy)."""
	A = laplacian_2d(40, np.float32)
	X = np.random.default_rng(1).standard_normal((1600, 32)).astype(np.float32)
	y2, y1 = _both(A, "exp", 20, X, dtype=np.float32, t=-0.2)
	assert np.max(np.abs(y2 - y1)) <= 2e-5 * np.max(np.abs(y1))
	rng = np.random.default_rng(5)
	B = rng.standard_normal((512, 512))
	D = B @ B.T / 512 + 0.3 * np.eye(512)
	Xd = rng.standard_normal((512, 32))
	y2, y1 = _both(D, "log", 24, Xd)
	assert np.max(np.abs(y2 - y1)) <= 1e-12 * np.max(np.abs(y1))
	ew, U = np.linalg.eigh(D)
	ref = (U * np.log(ew)) @ (U.T @ Xd)
	assert np.max(np.abs(y2 - ref)) < 1e-6 * np.max(np.abs(ref))


def test_auto_policy_goes_basis_free_when_the_basis_would_not_fit(golden, monkeypatch):
	"""`basis="auto"`: stored basis while 4 columns of it fit in the budget, the two-pass engine beyond (here the budget is
	shrunk instead of the problem grown); both give the same f(A) X."""
	import torch

	from primate_b200 import operators
	from primate_b200.operators import MatrixFunction

	A = _L30(golden)
	M = MatrixFunction(A, "exp", deg=30, orth=0, t=-0.3)
	assert not M._use_twopass(900) and M._basis_block_width(900) == 32
	X = torch.from_numpy(np.random.default_rng(0).standard_normal((900, 8))).cuda()
	y_stored = M.matmat_device(X).cpu().numpy()
	monkeypatch.setattr(operators, "BASIS_BUDGET", 4 * 30 * 900 * 8 - 1)
	assert M._use_twopass(900) and M._basis_block_width(900) == 32
	y_two = M.matmat_device(X).cpu().numpy()
	assert np.max(np.abs(y_two - y_stored)) <= 1e-12 * np.max(np.abs(y_stored))
	assert not MatrixFunction(A, "exp", deg=30, orth=3)._use_twopass(900)  # re-orthogonalisation needs the stored vectors
