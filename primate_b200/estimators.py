"""Estimators and composable stopping criteria (API mirror of src/primate/estimators.py:1-354).

Host-side and scalar-only, exactly as in the reference: the device hands over per-probe quadrature values (or a
pre-reduced (count, mean, M2) record, see `MeanEstimator.merge`) and the criteria decide when to stop.
"""
import operator
from typing import Callable, Iterable, Literal, Optional, Protocol, Sized, Union, runtime_checkable

import numpy as np
import scipy as sp
from numpy.typing import ArrayLike

from .stats import Covariance, Mean
from .typing import restrict_kwargs


def arr_summary(x: Union[None, float, np.ndarray]) -> str:
	"""Short printable form of a scalar / small array / long array (head ... tail)."""
	if x is None:
		return "None"
	x = np.atleast_1d(x)
	with np.printoptions(precision=2, suppress=True, threshold=3, floatmode="fixed"):
		if len(x) == 1:
			return f"{x.item():.3f}"
		if len(x) <= 3:
			return np.array2string(x, separator=",")
		head = np.array2string(x[:2], separator=",").strip("[]")
		tail = np.array2string(x[-1], separator=",").strip("[]")
		return "[" + head + ",...," + tail + "]"


@runtime_checkable
class Estimator(Sized, Protocol):
	"""Protocol of an updateable estimator: counts its samples, exposes the current `estimate` and last `delta`."""

	n_samples: int = 0
	values: Optional[list] = None
	delta: Union[float, np.ndarray] = np.inf

	def __len__(self) -> int:
		return self.n_samples

	def update(self, x: ArrayLike): ...

	@property
	def estimate(self) -> Union[float, np.ndarray]: ...


class ConvergenceCriterion:
	"""Lazily evaluated predicate on an estimator; composes with `|`, `&`, `~` (each composition is a new criterion
	that evaluates its parts on the estimator it is called with)."""

	def __init__(self, operation: Callable):
		assert callable(operation)
		self._operation = operation

	@staticmethod
	def _compose(op: Callable, *parts: "ConvergenceCriterion") -> "ConvergenceCriterion":
		return ConvergenceCriterion(lambda est: op(*(part(est) for part in parts)))

	def __or__(self, other: "ConvergenceCriterion"):
		return self._compose(operator.or_, self, other)

	def __and__(self, other: "ConvergenceCriterion"):
		return self._compose(operator.and_, self, other)

	def __invert__(self):
		return self._compose(operator.not_, self)

	def __call__(self, est: Estimator) -> bool:
		return self._operation(est)

	def message(self, est: Estimator) -> str:
		return "Composite convergence criterion"


class EstimatorResult:
	"""What `full=True` returns and what callbacks receive: the estimator, the stopping criterion, the current
	estimate, a status message, the number of iterations and a free-form `info` dict (unpacks in that order)."""

	_FIELDS = ("estimator", "criterion", "estimate", "message", "nit", "info")

	def __init__(self, estimator: Optional[Estimator] = None, criterion: Union[ConvergenceCriterion, str, None] = None,
	             estimate: Union[float, np.ndarray] = 0.0, message: str = "", nit: int = 0, info: Optional[dict] = None):  # fmt: skip
		self.estimator, self.criterion, self.estimate = estimator, criterion, estimate
		self.message, self.nit, self.info = message, nit, ({} if info is None else info)

	def __iter__(self) -> Iterable:
		return (getattr(self, name) for name in self._FIELDS)

	def __repr__(self) -> str:
		return "EstimatorResult(" + ", ".join(f"{name}={getattr(self, name)!r}" for name in self._FIELDS) + ")"


class MeanEstimator(Estimator):
	"""Sample-mean estimator, optionally tracking the covariance (needed by `ConfidenceCriterion`)."""

	n_samples: int = 0
	delta: Union[float, np.ndarray] = np.atleast_1d(np.inf)
	values: Optional[list] = []

	def __init__(self, dim: int = 1, covariance: bool = False, record: bool = False) -> None:
		self.n_samples = 0
		self.delta = np.full(shape=dim, fill_value=np.inf)
		self.values = [] if record else None
		if covariance:
			self._cov = Covariance(dim=dim)
		else:
			self._mean = Mean(dim=dim)

	def _acc(self):
		return self._cov if hasattr(self, "_cov") else self._mean

	@property
	def mean(self) -> Union[float, np.ndarray]:
		if hasattr(self, "_cov"):
			mu = np.atleast_1d(self._cov.mean())
			return mu.item() if len(mu) == 1 else np.ravel(mu)
		return self._mean()

	def update(self, x: ArrayLike) -> None:
		"""x of shape (m,) = m scalar samples; (m, d) = m samples of dimension d."""
		x = np.atleast_1d(x)
		x = x[:, None] if x.ndim == 1 else x
		acc = self._acc()
		before = acc.mu.copy()
		acc.update(x)
		self.delta = acc.mu - before
		self.n_samples += x.shape[0]
		if self.values is not None:
			self.values.extend(x)

	def merge(self, count: int, mean, M2=0.0) -> None:
		"""Folds a pre-reduced (count, mean, M2) record (device block statistics / another rank) in one step."""
		if count <= 0:
			return
		acc = self._acc()
		before = acc.mu.copy()
		if hasattr(self, "_cov"):
			acc.merge(count, mean, M2)
		else:
			acc.merge(count, mean)
		self.delta = acc.mu - before
		self.n_samples += int(count)

	@property
	def estimate(self) -> Union[float, np.ndarray]:
		return self.mean


class ControlVariableEstimator(MeanEstimator):
	"""Mean estimator corrected by control variates with known expectation `ecv`."""

	def __init__(self, ecv: Union[float, np.ndarray], alpha: Union[float, np.ndarray, None] = None, record: bool = False):
		ecv = np.atleast_1d(ecv).ravel()
		super().__init__(len(ecv), covariance=False, record=record)
		if alpha is not None:
			alpha = np.atleast_1d(alpha).ravel()
			assert len(ecv) == len(alpha), "Coefficients alpha must have same length as the control variables."
		self.alpha = alpha
		self.ecv = ecv
		self.cov = Covariance(dim=len(ecv) + 1)
		self._estimate_cor = alpha is None
		self.n_samples = 0
		self.delta = np.inf

	def __len__(self) -> int:
		return self.n_samples

	def update(self, samples: ArrayLike):
		samples = np.atleast_1d(samples)
		self.cov.update(samples)
		self.n_samples = self.cov.n
		C = self.cov(ddof=1)
		if self._estimate_cor:
			self.alpha = (C[0, 1] / C[1, 1]) if self.cov.dim == 2 else np.linalg.solve(C[1:, 1:], C[1:, 0])
		return self

	@property
	def estimate(self):
		if self.n_samples == 0:
			return np.nan
		return (self.cov.mu[0] - np.dot(self.alpha, self.cov.mu[1:] - self.ecv)).item()


class CountCriterion(ConvergenceCriterion):
	"""True once the estimator has seen at least `count` samples."""

	def __init__(self, count: int):
		self.count = count

	def __call__(self, est: Estimator) -> bool:
		return len(est) >= self.count

	def message(self, est: Estimator) -> str:
		return f"Est: {arr_summary(np.array(est.estimate))} (#S:{ len(est) })"


class ToleranceCriterion(ConvergenceCriterion):
	"""True once the last change of the estimate is below atol, or below rtol relative to the estimate."""

	def __init__(self, rtol: float = 0.01, atol: float = 1.49e-08, ord: Union[Literal["fro", "nuc"], float, None] = 2.0) -> None:
		self.rtol = rtol
		self.atol = atol
		self.ord = ord

	def __call__(self, est: Estimator) -> bool:
		if est.estimate is None:
			return False
		change = np.linalg.norm(est.delta, ord=self.ord)
		size = np.linalg.norm(np.atleast_1d(est.estimate), ord=self.ord)
		return bool(change < self.atol or change < self.rtol * size)

	def message(self, est: Estimator) -> str:
		msg = f"Est: {arr_summary(est.estimate)}"
		msg += f"(atol={self.atol:3f}, rtol={self.rtol:3f}, #S:{ len(est) })"
		if est.estimate is not None:
			change = np.linalg.norm(np.atleast_1d(est.delta), ord=self.ord)
			size = np.linalg.norm(np.atleast_1d(est.estimate), ord=self.ord)
			msg += f"\nnorm(it - est, {self.ord}) = {change:.3f}, norm(est, {self.ord}) = {size:.3f}"
		return msg


class ConfidenceCriterion(ConvergenceCriterion):
	"""CLT-based stop: margin of error of the sample mean <= atol, or relative standard error <= rtol."""

	def __init__(self, confidence: float = 0.95, atol: float = 0.00, rtol: float = 0.01) -> None:
		assert 0 < confidence and confidence < 1, "Confidence must be in (0, 1)"
		self.atol = 0.0 if atol is None else atol
		self.rtol = 0.0 if rtol is None else rtol
		self.z = 2 ** (1 / 2) * sp.special.erfinv(confidence)
		# NB the reference indexes this table by n_samples (estimators.py:254,286), i.e. df = n + 1; kept as is.
		self.t_scores = sp.stats.t.ppf((confidence + 1.0) / 2.0, df=np.arange(30) + 1)
		self.confidence = confidence

	def _error(self, est: MeanEstimator):
		if est.n_samples < 3:
			return (np.inf, np.inf)
		std_error = est._cov.covariance() ** (1 / 2) / np.sqrt(est.n_samples)
		score = self.t_scores[est.n_samples] if est.n_samples < 30 else self.z
		return (score * std_error, abs(std_error / est.estimate))

	def __call__(self, est: Estimator) -> bool:
		assert isinstance(est, MeanEstimator), "Must be a mean estimator"
		margin, rel = self._error(est)
		return margin <= self.atol or rel <= self.rtol

	def message(self, est: Estimator) -> str:
		margin, _ = self._error(est)
		return f"Est: {arr_summary(est.estimate)} +/- {margin:.3f} ({self.confidence*100:.0f}% CI, #S:{ len(est) })"


class KneeCriterion(ConvergenceCriterion):
	"""Kneedle detection on the recorded sample sequence (needs `record=True`)."""

	def __init__(self, S: float = 1.0) -> None:
		self.S = S

	def __call__(self, est: Estimator) -> bool:
		if est.values is None or len(est.values) < 3:
			return False
		seq = np.array(est.values).ravel()
		scaled = seq / np.arange(1, len(seq) + 1)
		curve = np.cumsum(np.abs(np.diff(scaled)))
		curve = (curve - curve.min()) / (curve.max() - curve.min())
		gap = curve - np.linspace(0, 1, len(curve))
		peak = gap[np.argmax(gap)]
		threshold = peak - (self.S / (len(curve) - 1))
		return peak > threshold and gap[-1] < threshold

	def message(self, est: Estimator) -> str:
		return f"Est: {arr_summary(est.estimate)} (#S:{ len(est) }, S={self.S:3f})"


CRITERIA = {
	"count": CountCriterion,
	"tolerance": ToleranceCriterion,
	"confidence": ConfidenceCriterion,
	"knee": KneeCriterion,
}


def convergence_criterion(criterion: Union[str, ConvergenceCriterion], **kwargs) -> ConvergenceCriterion:
	"""Builds a criterion from its name, passing only the keyword arguments it understands."""
	if isinstance(criterion, ConvergenceCriterion):
		return criterion
	assert isinstance(criterion, str) and criterion.lower() in CRITERIA.keys(), f"Invalid criterion {criterion}"
	cls = CRITERIA[criterion.lower()]
	cc = cls(**restrict_kwargs(cls, kwargs))
	assert isinstance(cc, ConvergenceCriterion), "`converge` must satisfy the ConvergenceEstimator protocol."
	return cc
