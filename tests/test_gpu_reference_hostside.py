"""Mirror of /root/reference/tests/test_tridiagonal.py:11-44 -- the tightest numeric pin on the path -- on
the device solvers: Lanczos with full re-orthogonalisation of a matrix with a prescribed spectrum, then the batched
implicit-QL eigensolver through `eigvalsh_tridiag`, `eigh_tridiag` (both method names) and the in-place `tqli`."""
import numpy as np
import pytest
from numpy.random import default_rng

pytestmark = pytest.mark.gpu


def test_tqli():
	"""tests/test_tridiagonal.py:11-23: in-place contract -- eigenvalues left in `d` (any order), `e` annihilated."""
	from primate_b200.lanczos import lanczos
	from primate_b200.random import symmetric
	from primate_b200.tridiag import tqli

	rng = default_rng(seed=1234)
	d = 50
	ew = np.sort(rng.uniform(size=d, low=1 / d, high=1))
	A = symmetric(d, seed=rng, pd=True, ew=ew)
	v = rng.uniform(size=d)
	a, b = lanczos(A, v0=v, deg=d, orth=d)
	d_, e_ = a.copy(), np.append([0], b)
	Z = np.empty((0, 0), dtype=A.dtype)
	tqli(d_, e_, Z, 30)
	assert np.allclose(np.sort(d_), ew)
	assert np.allclose(e_, 0.0)
	assert np.max(np.abs(np.sort(d_) - ew)) <= 1e-13
	# with eigenvectors: Z <- Z @ Y, columns orthonormal and T Y = Y diag(d) in the order tqli left them
	d2, e2, Z2 = a.copy(), np.append([0], b), np.eye(d)
	tqli(d2, e2, Z2, 30)
	T = np.diag(a) + np.diag(b, 1) + np.diag(b, -1)
	assert np.allclose(Z2.T @ Z2, np.eye(d), atol=1e-12)
	assert np.allclose(T @ Z2, Z2 * d2[None, :], atol=1e-12)
	# max_iter is honoured: with a single sweep per eigenvalue the iteration runs out and leaves off-diagonal mass
	d3, e3 = a.copy(), np.append([0], b)
	tqli(d3, e3, np.empty((0, 0)), 1)
	assert np.max(np.abs(e3)) > 1e-8 and abs(d3.sum() - a.sum()) < 1e-10  # similarity transforms keep the trace


@pytest.mark.parametrize("seed", [1234, 4756, 43, 102])
def test_tridiag(seed):
	"""tests/test_tridiagonal.py:26-44 at n = 150: eigenvalues vs the prescribed spectrum to 1e-14, orthonormal vectors."""
	from primate_b200.lanczos import lanczos
	from primate_b200.random import symmetric
	from primate_b200.tridiag import eigh_tridiag, eigvalsh_tridiag, tqli

	rng = default_rng(seed=seed)
	d = 150
	ew = np.sort(rng.uniform(size=d, low=1 / d, high=1))
	A = symmetric(d, seed=rng, pd=True, ew=ew)
	v = rng.uniform(size=d)
	a, b = lanczos(A, v0=v, deg=d, orth=d)
	for method in ["tqli", "mrrr"]:
		ew_test = np.sort(eigvalsh_tridiag(a, b, method=method))
		assert np.allclose(ew_test, ew), f"Eigenvalue test failed for method = {method}"
		assert np.max(np.abs(ew_test - ew)) <= 1e-14
	for method in ["tqli", "mrrr"]:
		ew_test, ev_test = eigh_tridiag(a, b, method=method)
		B = ev_test.T @ ev_test
		assert np.allclose(B - np.diag(B.diagonal()), 0.0)
		assert np.allclose(B.diagonal(), 1.0)
		assert np.allclose(np.sort(ew_test), ew)
	dd, ee = a.copy(), np.append([0], b)
	tqli(dd, ee, np.empty((0, 0)), 30)
	assert np.max(np.abs(np.sort(dd) - ew)) <= 1e-14
	# the device solver against LAPACK stemr on the same T (what the reference's default method calls, tridiag.py:10-11)
	from scipy.linalg import eigh_tridiagonal

	w_ref, Y_ref = eigh_tridiagonal(a, b)
	w, Y = eigh_tridiag(a, b)
	assert np.max(np.abs(w - w_ref)) <= 1e-14
	assert np.max(np.abs(np.square(Y[0, :]) - np.square(Y_ref[0, :]))) <= 5e-12  # the Golub-Welsch weights
