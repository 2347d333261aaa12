"""Lanczos tridiagonalisation front-end (API mirror of src/primate/lanczos.py:25-164).

`_lanczos.lanczos(A, v, deg, rtol, orth, alpha, beta, Q)` keeps the in-place contract of the reference's pybind11
module (src/primate/_lanczos.cpp:86-100) on host arrays -- it is the k = 1 case of the blocked device driver
(libslq: slq_lanczos).  Operators are uploaded once and cached per Python object, instead of being converted and
copied on every call.
"""
import weakref
from typing import Any, Optional, Union

import numpy as np
from scipy.sparse import sparray, spdiags
from scipy.sparse.linalg import LinearOperator

from . import engine
from .tridiag import eigh_tridiag, eigvalsh_tridiag

_OP_CACHE: dict = {}


def device_operator(A: Any, dtype=None) -> "engine.DeviceOperator":
	"""DeviceOperator for `A`, cached on the identity of the host object (and dtype)."""
	if isinstance(A, engine.DeviceOperator):
		return A
	dt = np.dtype(dtype) if dtype is not None else np.dtype(A.dtype if hasattr(A, "dtype") else np.float64)
	key = (id(A), dt.str)
	hit = _OP_CACHE.get(key)
	if hit is not None and hit[0]() is A:
		return hit[1]
	op = engine.DeviceOperator(A, dtype=dt)
	try:
		ref = weakref.ref(A, lambda _r, k=key: _OP_CACHE.pop(k, None))
	except TypeError:  # object without weakref support (e.g. np.matrix subclasses): do not cache
		return op
	if len(_OP_CACHE) >= 8:
		_OP_CACHE.pop(next(iter(_OP_CACHE)))
	_OP_CACHE[key] = (ref, op)
	return op


class _LanczosFFI:
	"""Stands where the reference's extension module `primate._lanczos` stands."""

	@staticmethod
	def lanczos(A, v: np.ndarray, deg: int, rtol: float, orth: int, alpha: np.ndarray, beta: np.ndarray, Q: np.ndarray) -> None:
		dt = np.dtype(v.dtype)
		if dt.type not in {np.float32, np.float64} or alpha.dtype != dt or beta.dtype != dt or Q.dtype != dt:
			raise TypeError("lanczos(): incompatible function arguments (all arrays must share a float32/float64 dtype)")
		op = device_operator(A, dt)
		ncv = int(Q.shape[1])
		X = engine.block_from_columns(np.asarray(v).reshape(-1, 1), dt)
		out = op.lanczos(X, int(deg), float(rtol), int(orth), ncv, want_basis=True)
		alpha[: deg + 1] = out["alpha"].cpu().numpy()[0]
		beta[: deg + 1] = out["beta"].cpu().numpy()[0]
		Q[:, :] = out["Q"][:, :, 0].t().cpu().numpy()


_lanczos = _LanczosFFI()


def lanczos(
	A: Union[np.ndarray, sparray, LinearOperator],
	v0: Optional[np.ndarray] = None,
	deg: Optional[int] = None,
	rtol: float = 1e-8,
	orth: int = 0,
	sparse_mat: bool = False,
	return_basis: bool = False,
	seed: Union[int, np.random.Generator, None] = None,
	dtype: Optional[np.dtype] = None,
	**kwargs: Any,
) -> tuple:
	"""Runs `deg` steps of the Lanczos method (Paige's A27 variant) on the symmetric operator `A` from `v0`.

	Returns the diagonal/off-diagonal `(a, b)` of the Jacobi matrix T = Q^T A Q (or T as a sparse matrix when
	`sparse_mat`), plus the Krylov basis `Q` when `return_basis`.  `orth` previous Lanczos vectors are re-projected
	out per step (0 = none, >= deg = full re-orthogonalisation).  Pre-allocated `alpha`, `beta`, `Q` may be passed
	as keyword arguments and are written in place.
	"""
	n: int = A.shape[0]
	deg: int = A.shape[1] if deg is None else min(deg, A.shape[1])
	f_dtype = (A @ np.zeros(A.shape[1])).dtype if not hasattr(A, "dtype") else A.dtype
	dt = dtype if dtype is not None else f_dtype
	assert deg > 0, "Number of steps must be positive!"
	assert np.dtype(f_dtype).type in {np.float32, np.float64}, "Only 32- or 64-bit floating point numbers are supported."

	orth: int = deg if orth < 0 or orth > deg else orth
	ncv: int = int(np.clip(orth, 2, deg)) if not (return_basis) else deg

	if v0 is None:
		rng = np.random.default_rng(seed)
		v0 = rng.uniform(size=A.shape[1], low=-1.0, high=+1.0).astype(dt)
	else:
		v0 = np.array(v0).astype(dt)
	assert len(v0) == A.shape[1], "Invalid starting vector; must match the number of columns of A."

	alpha = kwargs.get("alpha", np.zeros(deg + 1, dtype=f_dtype))
	beta = kwargs.get("beta", np.zeros(deg + 1, dtype=f_dtype))
	Q = kwargs.get("Q", np.zeros((n, ncv), dtype=f_dtype, order="F"))
	assert isinstance(alpha, np.ndarray) and len(alpha) == deg + 1 and alpha.dtype == f_dtype and alpha.flags["WRITEABLE"]
	assert isinstance(beta, np.ndarray) and len(beta) == deg + 1 and beta.dtype == f_dtype and beta.flags["WRITEABLE"]
	assert Q.ndim == 2 and Q.shape == (n, ncv) and Q.flags["F_CONTIGUOUS"] and Q.flags["WRITEABLE"] and Q.flags["OWNDATA"]

	_lanczos.lanczos(A, v0.astype(f_dtype, copy=False), deg, rtol, orth, alpha, beta, Q)

	if sparse_mat:
		T = spdiags(data=[np.roll(beta, -1), alpha, beta], diags=(-1, 0, +1), m=deg, n=deg)
		return T if not return_basis else (T, Q)
	a, b = alpha[:deg], beta[1:deg]
	return (a, b) if not return_basis else ((a, b), Q)


def rayleigh_ritz(A, deg: Optional[int] = None, return_eigenvectors: bool = False, method: str = "RRR", **kwargs) -> Union[np.ndarray, tuple]:
	"""Rayleigh-Ritz approximations of the eigenvalues of `A`: Lanczos up to degree `deg` (clipped to [2, n]) followed
	by the tridiagonal eigen-solver.  `return_eigenvectors` adds the Ritz vectors Y of T; `return_basis=True` (kwarg)
	adds the Lanczos basis Q.  No ghost-eigenvalue filtering is performed (as in the reference)."""
	n: int = A.shape[0]
	deg: int = A.shape[1] if deg is None else min(deg, A.shape[1])
	assert deg > 0, "Number of steps must be positive!"
	deg = int(np.clip(deg, 2, n))
	with_basis = kwargs.pop("return_basis", False)
	out = lanczos(A, deg=deg, return_basis=with_basis, **kwargs)
	(a, b), Q = out if with_basis else (out, None)
	if return_eigenvectors:
		rw, Y = eigh_tridiag(a, b)
		return (rw, Y) if not with_basis else (rw, Y, Q)
	rw = eigvalsh_tridiag(a, b)
	return rw if not with_basis else (rw, Q)
