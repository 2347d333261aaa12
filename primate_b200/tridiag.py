"""Symmetric tridiagonal eigen-solvers on the device (API mirror of src/primate/tridiag.py:7-62, tqli.py:15-90).

The reference calls LAPACK stemr through SciPy (method "mrrr") with an in-tree implicit-QL fallback ("tqli").  Here
every method name runs the same batched device solver (csrc/slq_quadrature.cu: one warp per Jacobi matrix,
implicit-shift QL); eigenvalues come back ascending (LAPACK order) and eigenvectors are only defined up to sign, as
with either reference method.
"""
import numpy as np

from . import _ffi, engine


def _prep(d: np.ndarray, e: np.ndarray):
	assert len(d) in {len(e) + 1, len(e)}, "Invalid diagonal/subdiagonal pair"
	d = np.asarray(d)
	e = np.append([0], e) if len(e) == (len(d) - 1) else np.asarray(e)
	dt = np.float32 if d.dtype == np.float32 else np.float64
	return np.ascontiguousarray(d, dtype=dt), np.ascontiguousarray(e, dtype=dt), dt


def _device_eigh(d: np.ndarray, e: np.ndarray, vectors: bool):
	"""d (m,), e (m,) with e[0] = 0  ->  ascending eigenvalues [and eigenvectors (m, m)] via the device QL solver."""
	torch = engine._torch()
	m = len(d)
	a = torch.from_numpy(d[None, :].copy()).cuda()
	b = torch.from_numpy(e[None, :].copy()).cuda()
	if not vectors:
		theta, _, _, info = engine.quadrature(a, b, m, fun_id=-1)
		return theta.cpu().numpy()[0].astype(d.dtype)
	_, theta, _, scratch, info = engine.funm_coeffs(a, b, m, fun_id=-1)
	scratch = scratch.cpu().numpy()
	Zt = scratch[: m * m].reshape(m, m)  # Zt[col, row]
	perm = scratch[m * m :].astype(np.int64)
	return theta.cpu().numpy()[0].astype(d.dtype), np.ascontiguousarray(Zt[perm, :].T).astype(d.dtype)


def _eigh_tridiag(d: np.ndarray, e: np.ndarray, Z: np.ndarray, method: str = "auto", maxiter: int = 30):
	want_vectors = np.prod(Z.shape) != 0
	return _device_eigh(d, e, want_vectors)


def eigh_tridiag(d: np.ndarray, e: np.ndarray, method: str = "auto", maxiter: int = 30):
	"""Eigenpairs (w, Y) of the symmetric tridiagonal matrix with diagonal `d` and subdiagonal `e`
	(`e` of length n with a leading zero, or n-1)."""
	assert method in {"tqli", "mrrr", "auto"}
	d, e, _ = _prep(d, e)
	return _device_eigh(d, e, True)


def eigvalsh_tridiag(d: np.ndarray, e: np.ndarray, method: str = "auto", maxiter: int = 30):
	"""Eigenvalues (ascending) of the symmetric tridiagonal matrix (d, e)."""
	assert method in {"tqli", "mrrr", "auto"}
	d, e, _ = _prep(d, e)
	return _device_eigh(d, e, False)


def tqli(d: np.ndarray, e: np.ndarray, Z: np.ndarray, max_iter: int = 30) -> None:
	"""In-place contract of the reference's Pythran kernel (tqli.py:15): on return `d` holds the eigenvalues and, when
	Z is m x m, its columns the eigenvectors of the input matrix times the incoming Z.  `e` has a leading zero."""
	dd, ee, _ = _prep(np.array(d, copy=True), np.array(e, copy=True))
	if np.prod(Z.shape) == 0:
		d[:] = _device_eigh(dd, ee, False)
	else:
		w, Y = _device_eigh(dd, ee, True)
		d[:] = w
		Z[:] = Z @ Y
	e[:] = 0
