"""Running mean / covariance (API mirror of src/primate/stats.py:7-113).

`Covariance.update` is the batched Chan-Golub-LeVeque merge the reference uses (stats.py:66-86); `merge` applies
the same rule to an already-reduced record (count, mean, M2), which is how device-side block statistics
(libslq: slq_stats_update) and other ranks' records (one NCCL all-gather per convergence check) enter the
estimator without moving samples to the host.
"""
from numbers import Number
from typing import Union

import numpy as np
import scipy as sp


def _as_samples(X, dim: int) -> np.ndarray:
	X = np.atleast_1d(X)
	X = X[:, None] if X.ndim == 1 else X
	assert X.shape[1] == dim, f"Expected shape (n, {dim}), got {X.shape}"
	return X


class Mean:
	"""Updateable sample mean of `dim`-dimensional observations."""

	def __init__(self, dim: int = 1):
		self.dim = dim
		self.n = 0
		self.mu = np.zeros(dim)

	def mean(self) -> Union[float, np.ndarray]:
		if self.n == 0:
			return np.nan
		return self.mu.item() if self.dim == 1 else self.mu

	__call__ = mean

	def update(self, X: Union[float, np.ndarray]) -> None:
		"""Folds a (batch, dim) array of new observations into the mean."""
		X = _as_samples(X, self.dim)
		m = X.shape[0]
		total = self.n + m
		self.mu += (m / total) * (X.mean(axis=0) - self.mu)
		self.n = total

	def merge(self, count: int, mean) -> None:
		"""Folds a pre-reduced record (count, mean) into the mean."""
		if count <= 0:
			return
		total = self.n + int(count)
		self.mu += (count / total) * (np.atleast_1d(mean) - self.mu)
		self.n = total


class Covariance(Mean):
	"""Updateable mean and (co)variance via Welford/Chan merging of batches."""

	def __init__(self, dim: int = 1):
		super().__init__(dim)
		self.S = np.zeros((dim, dim))

	def covariance(self, ddof: int = 1) -> Union[float, np.ndarray]:
		"""Covariance estimate S / (n - ddof); infinite while n <= ddof."""
		if (self.n - ddof) <= 0:
			return np.inf if self.dim else np.diag(np.inf, self.dim)
		cov = self.S / (self.n - ddof)
		return cov.item() if self.dim == 1 else cov

	__call__ = covariance

	def _shift_term(self, delta: np.ndarray):
		return delta[:, None] @ delta[None, :] if self.dim > 1 else (delta * delta)

	def update(self, X: Union[float, np.ndarray]) -> None:
		"""Folds a (batch, dim) array of new observations into mean and scatter matrix."""
		X = _as_samples(X, self.dim)
		m = X.shape[0]
		batch_mu = X.mean(axis=0)
		delta = batch_mu - self.mu
		total = self.n + m
		self.mu += (m / total) * delta
		R = X - batch_mu
		self.S += (R.T @ R) + (self.n * m / total) * self._shift_term(delta)
		self.n = total

	def merge(self, count: int, mean, M2=0.0) -> None:
		"""Folds a pre-reduced record (count, mean, M2 = sum of squared deviations) -- same rule as `update`."""
		if count <= 0:
			return
		count = int(count)
		delta = np.atleast_1d(np.asarray(mean, dtype=np.float64)) - self.mu
		total = self.n + count
		self.mu += (count / total) * delta
		self.S += np.reshape(np.asarray(M2, dtype=np.float64), self.S.shape) + (self.n * count / total) * self._shift_term(delta)
		self.n = total


def confidence_interval(a: np.ndarray, confidence: float = 0.95, sdist: str = "t") -> tuple:
	"""Confidence interval of the sample mean of `a` under a t or normal sampling distribution."""
	assert isinstance(confidence, Number) and confidence >= 0.0 and confidence <= 1.0, "Invalid confidence measure"
	center = np.mean(a)
	if sdist == "t":
		half = sp.stats.t.ppf((1 + confidence) / 2.0, len(a) - 1) * sp.stats.sem(a, ddof=1)
		return center - half, center + half
	if sdist == "normal":
		return sp.stats.norm.interval(confidence, loc=center, scale=np.std(a, ddof=1) / np.sqrt(len(a)))
	raise ValueError(f"Unknown sampling distribution '{sdist}'.")
