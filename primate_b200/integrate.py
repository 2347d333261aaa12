"""Gaussian quadrature rules of Jacobi matrices (API mirror of src/primate/integrate.py:9-76)."""
from typing import Optional

import numpy as np

from . import engine
from .fttr import fttr
from .tridiag import eigvalsh_tridiag


def quadrature(
	d: np.ndarray,
	e: np.ndarray,
	deg: Optional[int] = None,
	quad: str = "gw",
	nodes: Optional[np.ndarray] = None,
	weights: Optional[np.ndarray] = None,
	**kwargs,
) -> tuple:
	"""Nodes (Ritz values) and weights (squared first eigenvector components) of the degree-`deg` Gauss rule of the
	Jacobi matrix with diagonal `d` and off-diagonal `e` (length n with a leading zero, or n-1).

	`quad` is "gw"/"golub_welsch" (eigen-decomposition; here the batched device QL solver) or "fttr" (forward
	three-term recurrence).  If both `nodes` and `weights` are given they are filled in place as well.
	"""
	d = np.asarray(d)
	deg = len(d) if deg is None else int(min(deg, len(d)))
	e = np.append([0], e) if len(e) == (len(d) - 1) else np.asarray(e)
	assert len(d) == len(e) and np.isclose(e[0], 0.0), "Subdiagonal first element 'e[0]' must be close to zero"
	if quad in {"gw", "golub_welsch"}:
		torch = engine._torch()
		dt = np.float32 if d.dtype == np.float32 else np.float64
		a = torch.from_numpy(np.ascontiguousarray(d[:deg], dtype=dt)[None, :].copy()).cuda()
		b = torch.from_numpy(np.ascontiguousarray(e[:deg], dtype=dt)[None, :].copy()).cuda()
		th, ta, _, _ = engine.quadrature(a, b, deg, fun_id=-1)
		theta, tau = th.cpu().numpy()[0].astype(d.dtype, copy=False), ta.cpu().numpy()[0].astype(d.dtype, copy=False)
	elif quad == "fttr":
		theta = eigvalsh_tridiag(d, e, **kwargs)
		tau = np.zeros(len(theta), dtype=theta.dtype)
		fttr(theta, d, e, deg, tau)
	else:
		raise ValueError(f"Invalid quadrature method '{quad}' supplied")
	if nodes is not None and weights is not None:
		assert len(nodes) == deg and len(weights) == deg, "`nodes` and `weights` output arrays must be `deg` in length."
		np.copyto(nodes, theta)
		np.copyto(weights, tau)
	return theta, tau
