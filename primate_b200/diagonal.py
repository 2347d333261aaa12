"""Diagonal estimators (API mirror of src/primate/diagonal.py:11-138).  Every product with `A` is an `A @ block`,
which for a `MatrixFunction` is blocked device Lanczos + basis combination (operators.MatrixFunction._matmat)."""
from typing import Callable, Optional, Union

import numpy as np
import scipy as sp

from .estimators import ConvergenceCriterion, EstimatorResult, MeanEstimator, convergence_criterion
from .operators import MatrixFunction, is_valid_operator
from .random import isotropic


def diag(
	A: Union[sp.sparse.linalg.LinearOperator, np.ndarray],
	pdf: Union[str, Callable] = "rademacher",
	converge: Union[str, ConvergenceCriterion] = "tolerance",
	seed: Union[int, np.random.Generator, None] = None,
	full: bool = False,
	callback: Optional[Callable] = None,
	record: bool = False,
	**kwargs,
) -> Union[float, tuple]:
	r"""Girard-Hutchinson estimate of the diagonal of `A`: the running ratio $\sum_i v_i \odot A v_i / \sum_i v_i \odot v_i$,
	fed to a `MeanEstimator(dim=n)` once per probe and stopped by `converge` (diagonal.py:11-92)."""
	f_dtype = is_valid_operator(A)
	N: int = A.shape[0]
	rng = np.random.default_rng(seed)
	draw = isotropic(pdf=pdf, seed=rng)
	estimator = MeanEstimator(dim=N, covariance=False, record=record)
	converge = convergence_criterion(converge, **kwargs)
	if np.prod(A.shape) == 0:
		return 0.0 if not full else (0.0, EstimatorResult())

	verbose = full or callback is not None
	result = EstimatorResult(estimator, converge) if verbose else None
	numer, denom = np.zeros(N, dtype=f_dtype), np.zeros(N, dtype=f_dtype)
	while not converge(estimator):
		v = draw(size=N).astype(f_dtype)
		u = (A @ v).ravel()
		numer += u * v.ravel()
		denom += np.square(v.ravel())
		estimator.update(np.atleast_2d(numer / denom))
		if callback is not None:
			callback(result)
	if verbose:
		result.estimate = estimator.estimate
		return (estimator.estimate, result)
	return estimator.estimate


def xdiag(A: np.ndarray, m: Optional[int] = None, pdf: str = "sphere", seed: Union[int, np.random.Generator, None] = None):
	"""XDiag (Epperly et al.) estimate of the diagonal of `A` from `m / 2` block products with `A` and `A.T`
	(diagonal.py:99-138).  On a `MatrixFunction` (symmetric: A.T = A) the n x m panels N, A N, Q, A Q stay on the GPU and
	are reduced by libslq's panel kernels (primate_b200.panels.DevicePanels); only m x m matrices reach the host."""
	from . import panels
	from .trace import _panel_backend

	m = 2 * A.shape[0] if m is None else min(m + (m % 2), 2 * A.shape[0])
	n, m = A.shape[0], m // 2
	rng = np.random.default_rng(seed=seed)
	P, finish = _panel_backend(A, pdf, rng, A.dtype if hasattr(A, "dtype") and np.dtype(A.dtype).kind == "f" else np.float64)

	Nw = P.probes(0, m)
	finish(m)
	Y = P.product(Nw)
	Q, R = P.qr(Y)
	lo, hi = P.row_slice()  # with one process per GPU: this rank's rows of the row-local part below
	Nr = P.rows(Nw, lo, hi)
	dNY = P.rowdot(Nr, P.rows(Y, lo, hi))
	del Y
	Z = P.product(Q) if P.on_device else A.T @ Q

	## m x m quantities (host)
	T = P.gram(Z, Nw)
	R_inv = panels.upper_inverse(R)
	S = R_inv.T / np.linalg.norm(R_inv, axis=1)
	dST = np.einsum("ij,ij->j", S, T)

	## diagonals of the n x n products, one row-dot of two panels each: row-local, so every rank does its slice
	Qr, Zr = P.rows(Q, lo, hi), P.rows(Z, lo, hi)
	dQZ = P.rowdot(Qr, Zr)
	dQSSZ = P.rowdot(P.matmul(Qr, S), P.matmul(Zr, S))
	dNTQ = P.rowdot(Nr, P.matmul(Qr, T))
	dNQSST = P.rowdot(Nr, P.matmul(Qr, S * dST[np.newaxis, :]))
	d = dQZ + (-dQSSZ + dNY - dNTQ + dNQSST) / m
	return np.ravel(P.to_host(P.gather_rows(d)))
