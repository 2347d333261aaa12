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
		engine.warn_unconverged(info, "eigvalsh_tridiag")
		return theta.cpu().numpy()[0].astype(d.dtype)
	_, theta, _, scratch, info = engine.funm_coeffs(a, b, m, fun_id=-1)
	engine.warn_unconverged(info, "eigh_tridiag")
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
	"""In-place contract of the reference's Pythran kernel (tqli.py:15-90), on the device (libslq: slq_tqli): on return
	`d` holds the eigenvalues in the order the QL sweep left them (not sorted), `e` (leading zero on input) the
	negligible off-diagonal remainders and, when Z is m x m, Z <- Z @ (eigenvectors of T).  At most `max_iter` implicit
	QL sweeps are spent per eigenvalue; like the reference, running out is tolerated silently."""
	torch = engine._torch()
	m = len(d)
	assert len(e) == m, "`e` must have the length of `d` (leading zero)"
	dd = torch.from_numpy(np.ascontiguousarray(d, dtype=np.float64)[None, :].copy()).cuda()
	ee = torch.from_numpy(np.ascontiguousarray(e, dtype=np.float64)[None, :].copy()).cuda()
	want_vectors = np.prod(Z.shape) != 0
	Zt = torch.from_numpy(np.ascontiguousarray(np.asarray(Z, dtype=np.float64).T)[None].copy()).cuda() if want_vectors else None
	_ffi.check(_ffi.lib().slq_tqli(dd.data_ptr(), ee.data_ptr(), Zt.data_ptr() if want_vectors else None, 1, m, int(max_iter), None,
	                               engine._stream_ptr()))  # fmt: skip
	d[:] = dd.cpu().numpy()[0]
	e[:] = ee.cpu().numpy()[0]
	if want_vectors:
		Z[:] = Zt.cpu().numpy()[0].T
