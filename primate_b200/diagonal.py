"""Diagonal estimators (API mirror of src/primate/diagonal.py:11-138).  Every product with `A` is an `A @ block`,
which for a `MatrixFunction` is blocked device Lanczos + basis combination (operators.MatrixFunction._matmat)."""
from typing import Callable, Optional, Union

import numpy as np
import scipy as sp

from .estimators import ConvergenceCriterion, EstimatorResult, MeanEstimator, convergence_criterion
from .operators import MatrixFunction, is_valid_operator
from .random import ProbeStream, isotropic


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
	(diagonal.py:99-138)."""
	m = 2 * A.shape[0] if m is None else min(m + (m % 2), 2 * A.shape[0])
	n, m = A.shape[0], m // 2
	if isinstance(A, MatrixFunction) and isinstance(pdf, str):
		return _xdiag_device(A, n, m, pdf, np.random.default_rng(seed=seed))

	def dprod(X, Y):  # diag(X Y^T) of two (n, m) panels, as a column
		return np.einsum("ij,ij->i", X, Y)[:, np.newaxis]

	rng = np.random.default_rng(seed=seed)
	draw = isotropic(pdf=pdf, seed=rng)

	Nw = draw(size=(n, m))
	Y = A @ Nw
	Q, R = np.linalg.qr(Y, mode="reduced")
	dNY = dprod(Nw, Y)
	del Y

	Z = A.T @ Q
	T = Z.T @ Nw
	R_inv = np.linalg.inv(R)
	S = R_inv.T / np.linalg.norm(R_inv, axis=1)
	QS = Q @ S

	dQZ = dprod(Q, Z)
	dQSSZ = dprod(QS, Z @ S)
	dNTQ = dprod(Nw, Q @ T)
	dST = np.einsum("ij,ij->j", S, T)  # diag(S^T T)
	dNQSST = dprod(Nw, QS * dST[np.newaxis, :])

	d = dQZ + (-dQSSZ + dNY - dNTQ + dNQSST) / m
	return d.ravel()


def _xdiag_device(A: MatrixFunction, n: int, m: int, pdf: str, rng: np.random.Generator) -> np.ndarray:
	"""XDiag with every n x m panel resident on the GPU: the same formulas as above on the same probes (device generator
	for Rademacher probes, host generator + one upload otherwise), block products from the Lanczos engine (columns split
	over the ranks when the MatrixFunction has a `comm`), thin QR / small inverses from torch on the same device."""
	import torch

	from . import engine

	stream = ProbeStream(n, pdf, seed=rng, dtype=A.dtype)
	cols, off = [], 0
	for k in engine.block_sizes(m, 64):
		cols.append(stream.block(k, off))
		off += k
	Nw = torch.cat(cols, dim=1) if len(cols) > 1 else cols[0]
	stream.finish(m)
	del cols

	Y = A.matmat_device(Nw)
	Q, R = torch.linalg.qr(Y, mode="reduced")
	dNY = (Nw * Y).sum(dim=1)
	del Y

	Z = A.matmat_device(Q)  # A.T @ Q: a matrix function of a symmetric operator is its own transpose
	T = Z.T @ Nw
	R_inv = torch.linalg.inv(R)
	S = R_inv.T / torch.linalg.norm(R_inv, dim=1)
	QS = Q @ S

	dQZ = (Q * Z).sum(dim=1)
	dQSSZ = (QS * (Z @ S)).sum(dim=1)
	dNTQ = (Nw * (Q @ T)).sum(dim=1)
	dST = (S * T).sum(dim=0)
	dNQSST = (Nw * (QS * dST[None, :])).sum(dim=1)

	d = dQZ + (-dQSSZ + dNY - dNTQ + dNQSST) / m
	return d.double().cpu().numpy().ravel()
