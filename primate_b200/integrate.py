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
		th, ta, _, info = engine.quadrature(a, b, deg, fun_id=-1)
		engine.warn_unconverged(info, "quadrature")
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


def spectral_density(
	A,
	deg: int = 50,
	count: int = 32,
	points=200,
	sigma: Optional[float] = None,
	pdf: str = "rademacher",
	seed=None,
	orth: int = 0,
	full: bool = False,
):
	r"""Stochastic-Lanczos estimate of the spectral density of a symmetric operator (Lin, Saad & Yang 2016): the average
	over `count` probes of the Gauss quadrature measures $\sum_i \tau_i \delta(t - \theta_i)$ of degree `deg`, smoothed
	by a Gaussian of width `sigma` (default: 2 x the mean node spacing).  The reference only sketches this consumer of
	the (theta, tau) arrays (`from .integrate import spectral_density` is commented out in src/primate/__init__.py:10);
	here all `count` Lanczos runs and Golub-Welsch solves happen on the device and only the (count x deg) nodes and
	weights come back.  `points`: number of evaluation points across the Ritz range, or an array of them.
	Returns (t, density) -- density integrates to one -- and with `full=True` also the raw (theta, tau) arrays."""
	from .operators import MatrixFunction
	from .random import ProbeStream

	M = A if isinstance(A, MatrixFunction) else MatrixFunction(A, "identity", deg=deg, orth=orth, dtype=A.dtype if hasattr(A, "dtype") else np.float64)
	stream = ProbeStream(M.shape[0], pdf, seed=seed, dtype=M.dtype)
	theta, tau, _ = M.nodes_range(stream, 0, int(count))
	stream.finish(int(count))
	theta, tau = theta.cpu().numpy(), tau.cpu().numpy()
	lo, hi = float(theta.min()), float(theta.max())
	if sigma is None:
		sigma = 2.0 * (hi - lo if hi > lo else 1.0) / theta.shape[1]
	if np.isscalar(points):
		t = np.linspace(lo - 5.0 * sigma, hi + 5.0 * sigma, int(points))  # the smoothed measure keeps its unit mass on the grid
	else:
		t = np.asarray(points, dtype=np.float64)
	z = (t[:, None] - theta.reshape(1, -1)) / sigma
	dens = (np.exp(-0.5 * z * z) * tau.reshape(1, -1)).sum(axis=1) / (np.sqrt(2 * np.pi) * sigma * theta.shape[0])
	return (t, dens, theta, tau) if full else (t, dens)
