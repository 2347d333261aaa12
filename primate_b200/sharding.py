"""Probe sharding over the GPUs of one box: one process per GPU, operator replicated, probes split.

Probes are a pure function of (seed, probe index) (src/primate/random.py:74-78 fills an F-ordered array in memory
order from ONE PCG64 stream), so any rank can generate any probe by jump-ahead; nothing on the data path crosses
GPUs.  The only exchange is one tiny collective per convergence check:

* fixed-count runs (`CountCriterion`): every rank reduces its quadrature samples on the device to a
  (count, mean, M2) record (libslq `slq_stats_update`), ONE all-gather of 3 doubles per rank follows and every rank
  folds the records in rank order with the Chan rule of src/primate/stats.py:77-86 -- numerically the reference's own
  batch update, without the cancellation of a raw (sum, sum-of-squares) all-reduce;
* adaptive criteria / `record=True`: the k samples of each rank's block are all-gathered instead, and every rank
  replays the reference's per-sample updates in global probe order, so all ranks stop at the same probe.

Block drivers (`hutchpp`, `xdiag`, any `M @ block` on a `MatrixFunction` built with `comm=`): the COLUMNS of the
block are independent Lanczos runs too, so they are split the same way (`sharded_columns`): every rank runs its
contiguous column range and the tall-skinny result panels are all-gathered, after which the small replicated algebra
(QR, inverse of R, ...) sees bit-identical data on every rank.

The collectives go through `torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np


class Comm:
	"""Minimal view of a process group: rank, size and two gathers of float64 payloads."""

	def __init__(self, group=None):
		import torch.distributed as dist

		self._dist = dist
		self.group = group
		self.rank = dist.get_rank(group)
		self.size = dist.get_world_size(group)
		self.backend = dist.get_backend(group)

	def _device(self):
		import torch

		return torch.device("cuda", torch.cuda.current_device()) if self.backend == "nccl" else torch.device("cpu")

	def all_gather(self, x) -> np.ndarray:
		"""x: (m,) float64 torch tensor (device or host) or ndarray, same m on every rank -> (size, m) ndarray."""
		import torch

		t = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64))
		t = t.to(self._device(), dtype=torch.float64).contiguous().reshape(-1)
		out = torch.empty(self.size * t.numel(), dtype=torch.float64, device=t.device)
		self._dist.all_gather_into_tensor(out, t, group=self.group)
		return out.cpu().numpy().reshape(self.size, -1)

	def all_gather_tensor(self, t):
		"""t: tensor on the backend's device, same shape and dtype on every rank -> (size, *t.shape), no host staging."""
		import torch

		t = t.contiguous()
		out = torch.empty((self.size,) + tuple(t.shape), dtype=t.dtype, device=t.device)
		self._dist.all_gather_into_tensor(out, t, group=self.group)
		return out

	def all_reduce_sum(self, t):
		"""In-place sum of a tensor on the backend's device over the ranks (every rank gets the same bits)."""
		self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.group)
		return t

	def barrier(self):
		self._dist.barrier(group=self.group)


class SoloComm:
	"""The single-process stand-in (no torch.distributed needed)."""

	rank, size, backend = 0, 1, "none"

	def all_gather(self, x) -> np.ndarray:
		try:
			import torch

			if isinstance(x, torch.Tensor):
				return x.detach().to("cpu", dtype=torch.float64).numpy().reshape(1, -1)
		except ImportError:  # pragma: no cover
			pass
		return np.asarray(x, dtype=np.float64).reshape(1, -1)

	def barrier(self):
		return None


def resolve_comm(comm="auto"):
	"""`comm="auto"`: the default torch.distributed group when one is initialised with more than one rank."""
	if comm is None or comm is False:
		return SoloComm()
	if isinstance(comm, (Comm, SoloComm)):
		return comm
	if comm == "auto":
		try:
			import torch.distributed as dist
		except ImportError:  # pragma: no cover
			return SoloComm()
		if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
			return Comm(None)
		return SoloComm()
	return Comm(comm)  # a ProcessGroup


def probe_range(total: int, rank: int, size: int) -> tuple:
	"""Contiguous probe-index range [lo, hi) of `rank` out of `size` for `total` probes."""
	per = -(-int(total) // int(size))
	lo = min(rank * per, total)
	return lo, min(lo + per, total)


def merge_records(records: Sequence[Sequence[float]]) -> tuple:
	"""Chan merge of (count, mean, M2) records in the given order (stats.py:77-86 applied to reduced batches)."""
	n, mu, S = 0.0, 0.0, 0.0
	for cnt, mean, m2 in records:
		if cnt <= 0:
			continue
		tot = n + cnt
		delta = mean - mu
		mu += (cnt / tot) * delta
		S += m2 + (n * cnt / tot) * delta * delta
		n = tot
	return int(n), mu, S


def record_of(samples: np.ndarray) -> tuple:
	"""(count, mean, M2) of a host sample vector."""
	x = np.asarray(samples, dtype=np.float64).ravel()
	if x.size == 0:
		return (0.0, 0.0, 0.0)
	mu = x.mean()
	return (float(x.size), float(mu), float(np.sum((x - mu) ** 2)))


def round_blocks(done: int, k: int, rank: int, size: int) -> tuple:
	"""Adaptive rounds: in the round starting at global probe `done`, rank r owns probes [done + r*k, done + (r+1)*k)."""
	return done + rank * k, k


def sharded_columns(fn, X: np.ndarray, comm, chunk_bytes: int = 1 << 30) -> np.ndarray:
	"""Y = fn(X) column by column, with the columns of X split over the ranks of `comm` (block drivers).

	`fn` maps an (n, j) host block to an (r, j) host block, column-wise independent (f(A) X, or the quadratic forms of
	the columns with r = 1).  Every rank must call this with the same X.  Rank r computes the contiguous column range
	`probe_range(m, r, size)`; the panels are all-gathered (padded to the common width, in pieces of at most
	`chunk_bytes` per rank) and every rank returns the same full (r, m) array."""
	X = np.asarray(X)
	m = X.shape[1]
	if comm.size <= 1 or m < 2:
		return np.asarray(fn(X))
	lo, hi = probe_range(m, comm.rank, comm.size)
	per = -(-m // comm.size)
	mine = np.asarray(fn(X[:, lo:hi])) if hi > lo else None
	rows = comm.all_gather(np.array([0.0 if mine is None else float(mine.shape[0])])).max()
	rows = int(rows)
	dtype = mine.dtype if mine is not None else X.dtype
	out = np.empty((rows, m), dtype=dtype, order="F")
	step = max(1, int(chunk_bytes // max(1, per * 8)))  # rows per piece
	for r0 in range(0, rows, step):
		r1 = min(rows, r0 + step)
		pad = np.zeros((per, r1 - r0), dtype=np.float64)  # column-major panel piece, one row of `pad` per column
		if mine is not None:
			pad[: hi - lo] = mine[r0:r1].T
		allp = comm.all_gather(pad.reshape(-1)).reshape(comm.size, per, r1 - r0)
		for r in range(comm.size):
			a, b = probe_range(m, r, comm.size)
			if b > a:
				out[r0:r1, a:b] = allp[r, : b - a].T
	return out
