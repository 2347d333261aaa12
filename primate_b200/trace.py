"""Trace estimators (API mirror of src/primate/trace.py:34-315) driven by the device engine.

`hutch` keeps the reference's signature, defaults, stopping cadence and results, but the Monte-Carlo loop
(trace.py:107-116: draw one probe -> `A.quad(v)` -> `estimator.update` -> `converge(estimator)`) runs on blocks of
up to 32 probes generated IN GPU memory (bit-identical to `isotropic(pdf, seed)` drawn column by column), pushed
through the probe-blocked Lanczos + Golub-Welsch kernels, and only the per-probe quadrature values (k doubles per
block, or a 3-double (count, mean, M2) record per rank) come back.  The host then replays the reference's update /
check sequence on those scalars, so the estimate, the number of samples consumed and the final state of a
user-supplied Generator are the reference's.  With torch.distributed initialised (one process per GPU) the probes
are sharded over the ranks (primate_b200.sharding).
"""
from typing import Callable, Optional, Union

import numpy as np
from scipy.sparse.linalg import LinearOperator

from . import engine, sharding
from .estimators import (
	ConfidenceCriterion,
	ConvergenceCriterion,
	CountCriterion,
	EstimatorResult,
	MeanEstimator,
	convergence_criterion,
)
from . import panels
from .linalg import append_trinv
from .operators import MatrixFunction, is_valid_operator
from .random import ProbeStream, isotropic

KMAX = 64  # probes per device block (power of two <= 64): 64 amortises the CSR bytes best (C2: +6% over 32)
KSPEC = 32  # block speculated per rank by the adaptive per-probe criteria (wasted work is bounded by one block)


# ---- Monte-Carlo driver shared by the device paths (and by the CPU gloo tests, with an injected sampler) --------
def _to_host(y) -> np.ndarray:
	return y.detach().cpu().numpy().ravel() if hasattr(y, "detach") else np.asarray(y, dtype=np.float64).ravel()


def _concat(parts: list):
	if len(parts) == 1:
		return parts[0]
	if hasattr(parts[0], "detach"):
		import torch

		return torch.cat(parts)
	return np.concatenate(parts)


def _sample_range(sample_block: Callable, lo: int, hi: int, kmax: int):
	"""Quadrature samples of probes [lo, hi) as one (hi-lo,) array, computed in power-of-two blocks."""
	ranged = getattr(sample_block, "sample_range", None)
	if ranged is not None and hi > lo:
		return ranged(lo, hi - lo, kmax)  # device fast path: one batched Golub-Welsch launch for the whole range
	parts, first = [], lo
	for k in engine.block_sizes(hi - lo, kmax):
		parts.append(sample_block(first, k))
		first += k
	return _concat(parts) if parts else np.zeros(0)


def _pad(y, m: int):
	"""Pads a (<= m,) sample vector to length m (zeros) so that every rank contributes the same size."""
	have = int(y.shape[0])
	if have == m:
		return y
	if hasattr(y, "detach"):
		import torch

		return torch.cat([y, torch.zeros((m - have,), dtype=y.dtype, device=y.device)])
	return np.concatenate([np.asarray(y, dtype=np.float64), np.zeros(m - have)])


def _gather_ranges(y, total: int, comm) -> np.ndarray:
	"""All ranks' samples of a `total`-probe range in global probe order.  Every rank's row of the gathered table is
	cut to the length `probe_range` assigns it -- a genuine NaN sample (QL non-convergence, inv/sqrt of a bad Ritz
	value) is kept and reaches the estimator exactly as it would on one rank."""
	per = -(-int(total) // comm.size)
	table = comm.all_gather(_pad(y, per)).reshape(comm.size, per)
	parts = []
	for r in range(comm.size):
		lo, hi = sharding.probe_range(total, r, comm.size)
		parts.append(table[r, : hi - lo])
	return np.concatenate(parts)


def _run_fixed_count(sample_block, estimator, total: int, comm, kmax: int, stats_fn=None) -> int:
	"""`CountCriterion` with per-probe cadence: exactly `total` more probes; one collective at the end."""
	if total <= 0:
		return 0
	lo, hi = sharding.probe_range(total, comm.rank, comm.size)
	y = _sample_range(sample_block, lo, hi, kmax)
	if comm.size > 1 and estimator.values is None and stats_fn is not None:
		# (count, mean, M2) per rank, reduced on the device; ONE all-gather of 3 doubles per rank
		rec = stats_fn(y) if hi > lo else np.zeros(3)
		for cnt, mean, m2 in comm.all_gather(rec):
			estimator.merge(int(cnt), mean, m2)
		return total
	allv = _gather_ranges(y, total, comm) if comm.size > 1 else _to_host(y)
	assert allv.size == total
	if total > 1:
		estimator.update(allv[:-1])
	estimator.update(allv[-1:])  # the reference's last update is a single sample: keeps `estimator.delta` meaningful
	return total


def _run_per_probe(sample_block, estimator, converge, comm, kmax: int) -> int:
	"""Adaptive criterion, per-probe cadence (trace.py:114-116): speculate one block per rank, replay on the host."""
	consumed = 0
	while not converge(estimator):
		first, k = sharding.round_blocks(consumed, kmax, comm.rank, comm.size)
		vals = comm.all_gather(sample_block(first, k)).ravel()  # global probe order: rank-major == index order
		for v in vals:
			estimator.update(v)
			consumed += 1
			if converge(estimator):
				break
	return consumed


def _run_batched(sample_block, estimator, converge, batch: int, callback, result, comm, kmax: int) -> int:
	"""`full=True` / callback cadence (trace.py:107-110): `batch` probes per update, one check per batch."""
	consumed = 0
	while not converge(estimator):
		lo, hi = sharding.probe_range(batch, comm.rank, comm.size)
		y = _sample_range(sample_block, consumed + lo, consumed + hi, kmax)
		vals = _gather_ranges(y, batch, comm) if comm.size > 1 else _to_host(y)
		estimator.update(vals)
		consumed += batch
		callback(result)
	return consumed


def _device_sampler(A, stream: ProbeStream):
	"""probe range -> device quadrature samples for a MatrixFunction (SLQ) or an explicit matrix (x^T A x)."""
	if isinstance(A, MatrixFunction):
		def _slq(first: int, k: int):
			return A.quad_block(stream.block(k, first))

		_slq.sample_range = lambda first, count, kmax: A.quad_range(stream, first, count, kmax)
		return _slq
	from .lanczos import device_operator

	dev = device_operator(A, stream.dtype)

	def _quad(first: int, k: int):
		import torch

		X = stream.block(k, first)
		Y = dev.matmat(X)
		return (X.to(torch.float64) * Y.to(torch.float64)).sum(dim=0)

	return _quad


def _device_stats(y):
	"""(count, mean, M2) of device samples, reduced by libslq (slq_stats_update); 3 doubles stay on the device."""
	import torch

	state = torch.zeros(3, dtype=torch.float64, device=y.device)
	engine.stats_update(y.contiguous(), state)
	return state


def _on_device(A) -> bool:
	import scipy.sparse as sp

	return isinstance(A, MatrixFunction) or isinstance(A, np.ndarray) or sp.issparse(A)


def hutch(
	A: Union[LinearOperator, np.ndarray],
	batch: int = 32,
	pdf: Union[str, Callable] = "rademacher",
	converge: Union[str, ConvergenceCriterion] = "default",
	seed: Union[int, np.random.Generator, None] = None,
	full: bool = False,
	callback: Optional[Callable] = None,
	**kwargs,
) -> Union[float, tuple]:
	r"""Estimates the trace of a symmetric `A` (or of `f(A)` for a `MatrixFunction`) with the Girard-Hutchinson estimator.

	Parameters are the reference's (trace.py:34-43): `batch` probes per update when `full`/`callback` is used (one
	probe per update otherwise), `pdf` the isotropic distribution (or a sampler callable), `converge` a criterion name
	or object ("default" = 200 samples OR a 95% CLT margin below 1.0), `seed` an int or Generator, remaining keyword
	arguments parameterise the criterion.  Extra keyword (not in the reference): `comm` -- "auto" (default) shards the
	probes over the ranks of an initialised torch.distributed group, `None` keeps everything on this process.
	"""
	f_dtype = is_valid_operator(A)
	N: int = A.shape[0]
	comm = sharding.resolve_comm(kwargs.pop("comm", "auto"))

	rng = np.random.default_rng(seed)
	estimator = MeanEstimator(covariance=True, record=kwargs.pop("record", False))
	if converge == "default":
		converge = CountCriterion(count=200) | ConfidenceCriterion(confidence=0.95, atol=1.0, rtol=0.0)
	else:
		converge = convergence_criterion(converge, **kwargs)

	if np.prod(A.shape) == 0:
		return 0.0 if not full else (0.0, EstimatorResult(estimator, converge))

	batched = full or callback is not None
	result = EstimatorResult(estimator, converge)
	cb = (lambda x: x) if callback is None else callback

	if isinstance(pdf, str) and _on_device(A):
		stream = ProbeStream(N, pdf, seed=rng, dtype=f_dtype)
		sampler = _device_sampler(A, stream)
		if batched:
			consumed = _run_batched(sampler, estimator, converge, int(batch), cb, result, comm, KMAX)
		elif type(converge) is CountCriterion:
			consumed = _run_fixed_count(sampler, estimator, converge.count - len(estimator), comm, KMAX, _device_stats)
		else:
			consumed = _run_per_probe(sampler, estimator, converge, comm, KSPEC)
		stream.finish(consumed)
	else:
		# host operators / user samplers: the reference's loop, one host probe (or batch) at a time
		draw = isotropic(pdf=pdf, seed=rng) if isinstance(pdf, str) else pdf
		quad_form = (lambda v: A.quad(v)) if hasattr(A, "quad") else (lambda v: np.einsum("...i,...i->...", v.T, (A @ v).T))
		while not converge(estimator):
			v = draw(size=(N, batch if batched else 1)).astype(f_dtype)
			estimator.update(quad_form(v))
			if batched:
				cb(result)
	if batched:
		result.message = converge.message(estimator)
		result.estimate = estimator.estimate
		result.nit = len(estimator)
		return (estimator.estimate, result)
	return estimator.estimate


def _panel_backend(A, pdf, rng: np.random.Generator, f_dtype):
	"""(panel backend, finish) for a block driver: device panels + device probes for a `MatrixFunction` with a named
	pdf on a PCG64 generator, NumPy panels otherwise.  `finish(consumed)` leaves the user's Generator where the
	reference would (host samplers draw from it directly, so there it is a no-op)."""
	if isinstance(A, MatrixFunction) and isinstance(pdf, str):
		stream = ProbeStream(A.shape[0], pdf, seed=rng, dtype=f_dtype)
		if stream.on_device:
			return panels.DevicePanels(A, stream), stream.finish
	draw = isotropic(pdf=pdf, seed=rng) if isinstance(pdf, str) else pdf
	return panels.HostPanels(A, draw, f_dtype), (lambda consumed: None)


def hutchpp(
	A: Union[LinearOperator, np.ndarray],
	m: Optional[int] = None,
	batch: int = 32,
	mode: str = "reduced",
	pdf: Union[str, Callable] = "rademacher",
	seed: Union[int, np.random.Generator, None] = None,
	full: bool = False,
) -> Union[float, dict]:
	"""Hutch++ estimator (trace.py:119-182): exact trace on a sketched range + Hutchinson on the deflated remainder.

	`m` probes go to each of the two phases (default n // 3, bumped by `m % 3` as in the reference); `batch` is
	unused.  mode "full" evaluates the sketched part as diag(Q^T A Q), any other mode through `A.quad` when the operator
	has one.  On a `MatrixFunction` every n x m panel stays on the GPU (primate_b200.panels.DevicePanels)."""
	f_dtype = is_valid_operator(A)
	N: int = A.shape[0]
	rng = np.random.default_rng(seed)
	if np.prod(A.shape) == 0:
		return 0.0 if not full else (0.0, EstimatorResult())

	nb = (N // 3) if m is None else m
	nb += nb % 3
	P, finish = _panel_backend(A, pdf, rng, f_dtype)

	## range sketch: Q = orth(A W)
	Q = P.qr(P.product(P.probes(0, nb)))[0]
	rng_ests = P.coldot(P.product(Q), Q) if mode == "full" else P.quad(Q)

	## Hutchinson on the complement of range(Q): G <- (I - Q Q^T) G
	G = P.probes(nb, nb)
	P.axpy_matmul(G, Q, P.gram(Q, G))
	defl_ests = P.coldot(P.product(G), G)
	finish(2 * nb)

	estimate = float(np.sum(rng_ests) + np.sum(defl_ests) / nb)
	if not full:
		return estimate
	result = EstimatorResult()
	result.estimate = estimate
	result.nit = 2 * nb
	result.samples = np.concatenate([np.ravel(rng_ests), np.ravel(defl_ests)])
	return result.estimate, result


def _xtrace_samples(Wq: np.ndarray, H: np.ndarray, T: np.ndarray, R: np.ndarray, R_inv: np.ndarray) -> np.ndarray:
	"""The m leave-one-out XTrace samples (trace.py:185-227) from m x m matrices only: Wq = Q^T W, H = Q^T A Q,
	T = (A Q)^T W and the QR factor R of A W with its inverse.  The reference hands `_xtrace` its sampler, not the pdf's
	name, so the "sphere" rescaling branch (trace.py:209-214) never runs: scale = 1."""
	cd = lambda X, Y: np.einsum("ij,ij->j", X, Y)  # noqa: E731  (diag(X^T Y))
	S = R_inv.T / np.linalg.norm(R_inv, axis=1)  # unit columns: the direction each probe adds to the sketch
	sw = cd(S, Wq)
	shs = cd(S, H @ S)
	HW = H @ Wq
	resid = -cd(T, Wq) + cd(Wq, HW) + sw * cd(S, R - HW) + np.abs(sw) ** 2 * shs + cd(T - H.T @ Wq, S) * sw
	return np.trace(H) - shs + resid


def xtrace(
	A: Union[LinearOperator, np.ndarray],
	batch: int = 32,
	pdf: Union[str, Callable] = "sphere",
	converge: Union[str, ConvergenceCriterion] = "default",
	seed: Union[int, np.random.Generator, None] = None,
	full: bool = False,
	callback: Optional[Callable] = None,
	**kwargs,
) -> Union[float, tuple]:
	"""Epperly's exchangeable XTrace estimator (trace.py:233-315): grows a sketch batch by batch and averages the
	leave-one-out estimates.  Signature and quirks are the reference's: a `CountCriterion(n)` is always in force and a
	user `converge` is OR-ed onto that same object (trace.py:271-276), so the loop runs until the sketch has n columns.

	The sketch is appended to block-wise (two Gram-Schmidt sweeps against Q + a thin QR of the remainder: the batch form
	of the reference's `qr_insert` per column), and the samples are formed from m x m matrices; on a `MatrixFunction`
	the n x m panels W, Q, A Q live on the GPU."""
	assert batch >= 1, "Batch size must be positive."
	n = A.shape[0]
	callback = (lambda result: ...) if not callable(callback) else callback
	record = kwargs.pop("record", False)
	estimator = MeanEstimator(record=record)

	if converge == "default":
		converge = CountCriterion(count=n)
	else:
		converge = CountCriterion(count=n)
		converge |= convergence_criterion(converge, **kwargs)
	assert isinstance(converge, ConvergenceCriterion)

	result = EstimatorResult()
	rng = np.random.default_rng(seed)
	f_dtype = A.dtype if hasattr(A, "dtype") else np.float64
	P, finish = _panel_backend(A, pdf, rng, np.dtype(f_dtype) if np.dtype(f_dtype).kind == "f" else np.float64)
	W, Q, Z = P.empty(n), P.empty(n), P.empty(n)  # probes, orth(A W), A Q: n columns at most
	R, R_inv, m = np.zeros((0, 0)), np.zeros((0, 0)), 0
	while not converge(estimator):
		ns = min(A.shape[1] - m, int(batch))
		Nw = P.probes(m, ns)
		Qn, C, Rn = panels.block_orthogonalize(P, P.cols(Q, 0, m) if m else None, P.product(Nw))
		R_inv = append_trinv(R_inv, C if C is not None else np.zeros((0, ns)), panels.upper_inverse(Rn))
		R = np.block([[R, C if C is not None else np.zeros((0, ns))], [np.zeros((ns, m)), Rn]])
		P.assign(W, m, Nw)
		P.assign(Q, m, Qn)
		P.assign(Z, m, P.product(Qn))
		m += ns
		Wm, Qm, Zm = P.cols(W, 0, m), P.cols(Q, 0, m), P.cols(Z, 0, m)
		t_samples = _xtrace_samples(P.gram(Qm, Wm), P.gram(Qm, Zm), P.gram(Zm, Wm), R, R_inv)
		estimator = MeanEstimator(record=record)  # XTrace re-derives all m samples from the grown sketch
		estimator.update(np.ravel(t_samples))
		callback(result)
	finish(m)

	result.estimator = estimator
	result.estimate = estimator.estimate
	result.criterion = converge
	return (result.estimate, result) if full else result.estimate
