"""Spectral functions by name (API mirror of src/primate/special.py:7-107).

The NumPy callables returned here are what `MatrixFunction.fun` holds (the reference's contract); the engine maps
the built-in *names* onto device-side evaluations (csrc/slq_quadrature.cu: eval_fun) so that hutch() never moves
theta/tau to the host for them.  Arbitrary Python callables are evaluated on the host on the tiny (k, deg) node
array.
"""
from typing import Any, Callable, Optional, Union

import numpy as np

_BUILTIN_MATRIX_FUNCTIONS = ["identity", "abs", "sqrt", "log", "inv", "exp", "smoothstep", "numrank"]


def softsign(x: Optional[np.ndarray] = None, q: int = 1) -> Union[Callable, np.ndarray]:
	"""Polynomial soft sign  sum_{i<=q} x (1-x^2)^i c_i,  c_0 = 1, c_i = c_{i-1} (2i-1)/(2i), on clip(x, -1, 1)."""
	powers = np.arange(q + 1)
	coefs = np.ones(q + 1)
	for i in range(1, q + 1):
		coefs[i] = coefs[i - 1] * (2 * i - 1) / (2 * i)

	def _sign(t: Union[float, np.ndarray]) -> np.ndarray:
		t = np.atleast_2d(np.clip(t, -1.0, +1.0)).T
		return np.ravel(np.sum(t * (1 - t**2) ** powers * coefs, axis=1))

	return _sign if x is None else _sign(x)


def smoothstep(x: Optional[np.ndarray] = None, a: float = 0.0, b: float = 1.0, deg: int = 3) -> Union[Callable, np.ndarray]:
	"""Cubic Hermite step: 0 below a, 1 above b, 3y^2 - 2y^3 in between (y = (x-a)/(b-a))."""
	assert (deg % 2) == 1, "Degree must be odd"
	width: float = (b - a) if a != b else 1.0

	def _smoothstep(t):
		y = np.clip((t - a) / width, 0.0, 1.0)
		return 3 * y**2 - 2 * y**3

	return _smoothstep if x is None else _smoothstep(x)


def identity(x: Any) -> Any:
	return x


def exp(x: Optional[np.ndarray] = None, t: float = 1.0) -> Union[Callable, np.ndarray]:
	def _exp(u):
		return np.exp(t * u)

	return _exp if x is None else _exp(x)


def step(x: Optional[np.ndarray] = None, c: float = 0.0, nonnegative: bool = False) -> Union[Callable, np.ndarray]:
	def _step(u):
		u = np.abs(u) if nonnegative else u
		return np.where(u < c, 0.0, 1.0)

	return _step if x is None else _step(x)


def param_callable(fun: Union[str, Callable, None], **kwargs) -> Callable:
	"""Name -> parameterised callable; same names, defaults and assertion behaviour as special.py:78-107."""
	if isinstance(fun, str):
		assert fun in _BUILTIN_MATRIX_FUNCTIONS, "If given as a string, matrix_function be one of the builtin functions."
	if fun is None or fun == "identity":
		return identity
	if callable(fun):
		return fun
	if fun == "abs":
		return np.abs
	if fun == "sqrt":
		return np.sqrt
	if fun == "log":
		return lambda x: np.log(np.maximum(x, np.finfo(np.float64).eps))
	if fun == "inv":
		return np.reciprocal
	if fun == "exp":
		return exp(t=kwargs.pop("t", 1.0))
	if fun == "smoothstep":
		return smoothstep(a=kwargs.pop("a", 0.0), b=kwargs.pop("b", 1.0))
	if fun == "softsign":
		return softsign(q=kwargs.pop("q", 10))
	if fun == "numrank":
		return step(c=kwargs.pop("threshold", 0.000001), nonnegative=True)
	raise ValueError(f"Unknown function: {fun}.")
