"""Mirrors of the reference's own host-side test modules, run against primate_b200's mirrors of those modules
(CPU only: estimators, criteria, running statistics and the probe generators are scalar host code on this path).

  /root/reference/tests/test_estimators.py:17-157   MeanEstimator, ControlVariableEstimator, Count / Tolerance /
                                                     Confidence / Knee criteria, `&` `|` `~` composition, the factory
  /root/reference/tests/test_stats.py:5-31          Covariance vs np.var / np.cov, confidence_interval
  /root/reference/tests/test_random.py:6-98         isotropic moments and supports, column-wise == block draws,
                                                     Isotropic(threads=1) == isotropic, haar, symmetric
Same seeds, same thresholds as the reference's tests."""
from typing import Callable

import numpy as np


# ---- tests/test_estimators.py ---------------------------------------------------------------------------------------
def test_MeanEstimator():
	from primate_b200.estimators import MeanEstimator

	rng = np.random.default_rng(1234)
	mu, samples = MeanEstimator(), []
	for _ in range(25):
		samples.extend(rng.normal(size=10))
		mu.update(samples[-10:])
	assert np.allclose(np.mean(samples), mu.mean)
	assert isinstance(mu.estimate, float)

	rng = np.random.default_rng(1234)
	mu, samples = MeanEstimator(dim=2), []
	for _ in range(25):
		samples.extend(rng.normal(size=(10, 2)))
		mu.update(samples[-10:])
	assert np.allclose(np.mean(samples, axis=0), mu.mean)

	rng = np.random.default_rng(1234)
	mu, samples = MeanEstimator(dim=2, covariance=True), []
	for _ in range(25):
		samples.extend(rng.normal(size=(10, 2)))
		mu.update(samples[-10:])
	assert np.allclose(np.mean(samples, axis=0), mu.mean)
	assert np.allclose(np.cov(samples, rowvar=False), mu._cov.covariance())


def test_MeanEstimator_merge_equals_update():
	"""The pre-reduced (count, mean, M2) entry the device / other ranks use is the batch update in one step."""
	from primate_b200.estimators import MeanEstimator

	rng = np.random.default_rng(7)
	x = rng.normal(1e6, 1e3, size=(9, 32))
	a, b = MeanEstimator(covariance=True), MeanEstimator(covariance=True)
	for row in x:
		a.update(row)
		b.merge(len(row), row.mean(), np.sum((row - row.mean()) ** 2))
		assert len(a) == len(b) and np.isclose(a.estimate, b.estimate, rtol=1e-15)
		assert np.allclose(a.delta, b.delta, rtol=1e-9, atol=1e-9)
	assert np.isclose(a._cov.covariance(), b._cov.covariance(), rtol=1e-12)
	assert np.isclose(a._cov.covariance(), np.var(x, ddof=1), rtol=1e-10)


def test_ControlVariableEstimator():
	from primate_b200.estimators import ControlVariableEstimator, MeanEstimator

	rng = np.random.default_rng(1235)
	a = np.array([1, 2, 3, 1, 2])
	h = lambda x: np.min([x[0] + x[3], x[0] + x[2] + x[4], x[1] + x[2] + x[3], x[1] + x[4]])  # noqa: E731
	mu = 1339 / 1440
	mu_cv = 15 / 16
	h_cv = lambda x: np.min([x[0] + x[3], x[1] + x[4]])  # noqa: E731
	n_efficient = 0
	for _ in range(150):
		U = rng.uniform(low=0, high=1, size=(250, 5))
		y = np.apply_along_axis(h, 1, U * a)
		y_cv = np.apply_along_axis(h_cv, 1, U * a)
		est1, est2 = MeanEstimator(), ControlVariableEstimator(mu_cv)
		est1.update(y)
		est2.update(np.c_[y, y_cv])
		n_efficient += np.linalg.norm(est2.estimate - mu) <= np.linalg.norm(est1.estimate - mu)
	assert n_efficient >= 120
	# fixed coefficients, several control variates, incremental updates
	est = ControlVariableEstimator([mu_cv, 0.5], alpha=[0.3, -0.1])
	U = rng.uniform(size=(64, 5))
	y, y1, y2 = np.apply_along_axis(h, 1, U * a), np.apply_along_axis(h_cv, 1, U * a), U[:, 0]
	est.update(np.c_[y, y1, y2][:40])
	est.update(np.c_[y, y1, y2][40:])
	want = y.mean() - 0.3 * (y1.mean() - mu_cv) + 0.1 * (y2.mean() - 0.5)
	assert len(est) == 64 and np.isclose(est.estimate, want, rtol=1e-12)


def test_CountCriterion():
	from primate_b200.estimators import CountCriterion, MeanEstimator

	rng = np.random.default_rng(1234)
	mu, cc = MeanEstimator(), CountCriterion(10)
	assert not cc(mu)
	for _ in range(9):
		mu.update(rng.uniform(size=1, low=-1, high=+1).item())
		assert not cc(mu)
	mu.update(rng.uniform(size=1, low=-1, high=+1).item())
	assert len(mu) == 10 and cc(mu)
	assert "#S:10" in cc.message(mu)


def test_ToleranceCriterion():
	from primate_b200.estimators import MeanEstimator, ToleranceCriterion

	rng = np.random.default_rng(1234)
	mu = MeanEstimator(15)
	cc = ToleranceCriterion(atol=0, rtol=0.10, ord=1)
	while not cc(mu):
		mu.update(rng.uniform(size=(1, 15), low=-1, high=+1))
	error = np.linalg.norm(mu.delta, ord=1)
	assert error < (np.linalg.norm(mu.estimate, ord=1) * 0.10)
	assert (np.linalg.norm(mu.estimate, ord=1) * 0.10) <= 0.20
	assert "rtol=0.1" in cc.message(mu)


def test_ConfidenceCriterion():
	from primate_b200.estimators import ConfidenceCriterion, MeanEstimator

	rng = np.random.default_rng(1234)
	mu, containing_intervals = 5.0, 0
	for _ in range(1500):
		atol = 0.50
		est = MeanEstimator(covariance=True)
		cc = ConfidenceCriterion(confidence=0.95, atol=atol, rtol=0.0)
		while not cc(est):
			est.update(rng.normal(size=5, loc=mu, scale=1 / 2))
		containing_intervals += np.abs(mu - est.estimate) <= atol
	assert abs((containing_intervals / 1500) - cc.confidence) < (100 / 1500)
	assert "95% CI" in cc.message(est)


def test_KneeCriterion():
	from primate_b200.estimators import KneeCriterion, MeanEstimator

	rng = np.random.default_rng(1234)
	mu, cc = MeanEstimator(record=True), KneeCriterion(S=1.0)
	assert not cc(mu)
	while not cc(mu):
		mu.update(rng.normal(size=1, loc=0, scale=1))
	assert cc(mu)
	assert np.abs(mu.delta) <= 0.15
	assert not KneeCriterion()(MeanEstimator())  # nothing recorded: never converges


def test_CriterionComposability():
	from primate_b200.estimators import ConfidenceCriterion, CountCriterion, MeanEstimator

	rng = np.random.default_rng(1234)
	mu = MeanEstimator(covariance=True)
	cc1, cc2 = CountCriterion(200), ConfidenceCriterion(confidence=0.95, atol=0.50, rtol=0.0)
	assert not cc1(mu) and not cc2(mu)
	cc = cc1 & cc2
	assert isinstance(cc, Callable) and cc(mu) is False
	while not (cc1(mu) and cc2(mu)):
		assert not cc(mu)
		mu.update(rng.uniform(size=1, low=-1, high=+1).item())
	assert cc1(mu) and cc2(mu) and cc(mu)
	assert (~cc)(mu) is False and (~~cc)(mu) is True
	assert cc.message(mu) == "Composite convergence criterion"

	mu = MeanEstimator(covariance=True)
	cc = cc1 | cc2
	assert isinstance(cc, Callable) and cc(mu) is False
	while not (cc1(mu) or cc2(mu)):
		assert not cc(mu)
		mu.update(rng.uniform(size=1, low=-1, high=+1).item())
	assert (cc1(mu) or cc2(mu)) and cc(mu)
	assert (~cc1)(MeanEstimator()) is True


def test_converge_param():
	from primate_b200.estimators import ConvergenceCriterion, KneeCriterion, convergence_criterion

	assert isinstance(convergence_criterion("count", count=10), ConvergenceCriterion)
	assert isinstance(convergence_criterion("tolerance", atol=0.0, rtol=0.0), ConvergenceCriterion)
	assert isinstance(convergence_criterion("confidence", confidence=0.95), ConvergenceCriterion)
	cc = convergence_criterion("knee", S=10.0, count=3)  # foreign keyword arguments are dropped (typing.restrict_kwargs)
	assert isinstance(cc, KneeCriterion) and cc.S == 10.0
	assert convergence_criterion(cc) is cc


# ---- tests/test_stats.py ------------------------------------------------------------------------------------------------
def test_Covariance():
	from primate_b200.stats import Covariance

	rng = np.random.default_rng(1234)
	C, samples = Covariance(dim=1), []
	for _ in range(25):
		samples.extend(rng.normal(size=10))
		C.update(samples[-10:])
		assert np.isclose(np.var(samples, ddof=1), C.covariance())
		assert np.isclose(np.mean(samples), C.mean())
		assert len(samples) == C.n
	C = Covariance(dim=2)
	samples.clear()
	for _ in range(25):
		samples.extend(rng.normal(size=(10, 2)))
		C.update(samples[-10:])
		assert np.allclose(np.cov(samples, rowvar=False, ddof=1), C.covariance())
		assert np.allclose(np.mean(samples, axis=0), C.mean())
		assert len(samples) == C.n


def test_confidence_interval():
	from primate_b200.stats import confidence_interval

	rng = np.random.default_rng(1234)
	samples = rng.normal(size=1500, loc=0, scale=1 / 2)
	ci_normal = confidence_interval(samples, confidence=0.95, sdist="normal")
	ci_tdist = confidence_interval(samples, confidence=0.95, sdist="t")
	assert np.max(np.abs(np.array(ci_normal) - np.array(ci_tdist))) <= 1e-4
	assert ci_tdist[0] < samples.mean() < ci_tdist[1]


# ---- tests/test_random.py -----------------------------------------------------------------------------------------------
def test_isotropic():
	from scipy.stats import normaltest

	from primate_b200.random import isotropic

	rng = np.random.default_rng(seed=1235)
	for method in ["rademacher", "sphere", "normal"]:
		S = isotropic(size=(5, 1500), pdf=method, seed=rng)
		ES = sum([np.outer(s, s) for s in S.T]) / S.shape[1]
		assert np.max(np.abs(ES - np.eye(S.shape[0]))) <= 0.15
		if method == "rademacher":
			assert list(np.unique(S.ravel())) == [-1, +1]
		elif method == "sphere":
			assert np.allclose(np.linalg.norm(S, axis=0), np.sqrt(S.shape[0]))
		elif method == "normal":
			assert normaltest(S.ravel()).pvalue >= 0.05
	assert np.allclose(isotropic(size=(150, 5), seed=1234), isotropic(size=(150, 5), seed=1234))


def test_iso_order():
	"""Column-by-column draws from one generator == one block draw (the fact probe sharding rests on)."""
	from primate_b200.random import isotropic, symmetric

	rng = np.random.default_rng(seed=1234)
	A = symmetric(20, ew=rng.uniform(size=20, low=0.0, high=1.0))
	rng = np.random.default_rng(seed=1234)
	qe = [v @ A @ v for v in (np.ravel(isotropic(20, pdf="sphere", seed=rng)) for _ in range(150))]
	rng = np.random.default_rng(seed=1234)
	V = isotropic((A.shape[1], 150), pdf="sphere", seed=rng)
	assert np.allclose(np.ravel(qe), np.diag(V.T @ A @ V))
	for pdf in ("rademacher", "normal"):
		g1, g2 = np.random.default_rng(5), np.random.default_rng(5)
		cols = np.concatenate([isotropic((33, 1), pdf=pdf, seed=g1) for _ in range(7)], axis=1)
		assert np.array_equal(cols, isotropic((33, 7), pdf=pdf, seed=g2))


def test_Isotropic():
	from primate_b200.random import Isotropic, isotropic

	Iso = Isotropic(size=(50, 5), pdf="sphere", seed=1234, threads=1)
	Iso.fill()
	assert np.allclose(Iso.values, isotropic((50, 5), "sphere", seed=1234))
	Iso = Isotropic(size=(50, 1500), pdf="normal", seed=1234, threads=4)  # spawned sub-streams: a different, valid stream
	Iso.fill()
	assert Iso.values.shape == (50, 1500) and abs(Iso.values.std() - 1) < 0.02 and not np.any(Iso.values == 0)


def test_haar():
	from primate_b200.random import haar

	rng = np.random.default_rng(1234)
	assert np.allclose(haar(25, ew=np.ones(25), seed=rng), np.eye(25))
	A = haar(25, seed=rng)
	assert not np.all(A == A.T)


def test_symmetric():
	from primate_b200.random import symmetric

	rng = np.random.default_rng(1234)
	ew = rng.uniform(size=25)
	A = symmetric(25, ew=ew, seed=rng)
	assert np.allclose(A, A.T)
	assert np.allclose(np.sort(ew), np.sort(np.linalg.eigvalsh(A)))
