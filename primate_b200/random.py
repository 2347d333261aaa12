"""Isotropic probe vectors (API mirror of src/primate/random.py:11-201).

`isotropic` keeps the reference's host contract (an F-ordered float64 (N, k) array filled in memory order from one
NumPy PCG64 stream, random.py:22-29,74-78).  The estimators do not call it on the hot path: they use
`ProbeStream`, which produces the SAME probes directly in GPU memory -- Rademacher probes by jump-ahead into the
PCG64 stream on the device (csrc/slq_probes.cu; probe i = stream elements [i*N, (i+1)*N), sign = top bit of the raw
output), normal/sphere probes by the host generator plus one upload (NumPy's ziggurat consumes a data-dependent
number of raw draws, so the stream position of a normal variate is inherently sequential).  Either way
the user's Generator ends up advanced exactly as the reference would leave it.
"""
import concurrent.futures
import multiprocessing
from typing import Callable, Optional, Union

import numpy as np
import scipy as sp

_ISO_DISTRIBUTIONS = {
	"rademacher": "rademacher",
	"normal": "normal",
	"sphere": "sphere",
	"signs": "rademacher",
	"gaussian": "normal",
}


def _rademacher_inplace(rng: np.random.Generator, out: np.ndarray) -> None:
	# u in [0,1) -> floor(2u) in {0,1} -> {-1,+1}; equivalently the top bit of the raw 64-bit draw
	rng.random(out=out)
	np.multiply(out, 2, out=out)
	np.floor(out, out=out)
	np.multiply(out, 2, out=out)
	np.subtract(out, 1, out=out)


def _normal_inplace(rng: np.random.Generator, out: np.ndarray) -> None:
	rng.standard_normal(out=out, dtype=out.dtype)


def _sphere_inplace(rng: np.random.Generator, out: np.ndarray) -> None:
	rng.standard_normal(out=out, dtype=out.dtype)
	col_norms = np.sqrt(np.sum(out**2, axis=0, keepdims=True))
	np.divide(out, col_norms, out=out)
	np.multiply(out, np.sqrt(out.shape[0]), out=out)


_ISO_FUNCS = {"rademacher": _rademacher_inplace, "sphere": _sphere_inplace, "normal": _normal_inplace}


def isotropic(
	size: Union[int, tuple, None] = None,
	pdf: str = "rademacher",
	seed: Union[int, np.random.Generator, None] = None,
	out: Optional[np.ndarray] = None,
) -> Union[None, np.ndarray, Callable]:
	"""Random vectors from an isotropic distribution ("rademacher"/"signs", "normal"/"gaussian", "sphere").

	With `out` the array is filled in place; with `size=None` a sampler closure over one generator is returned.
	"""
	assert pdf in _ISO_DISTRIBUTIONS.keys(), f"Invalid distribution '{pdf}' supplied."
	fill = _ISO_FUNCS[_ISO_DISTRIBUTIONS[pdf]]
	rng = np.random.default_rng(seed)
	if out is not None:
		assert isinstance(out, np.ndarray)
		fill(rng, out)
		return None

	def _isotropic(size: Union[int, tuple]):
		shape = (size, 1) if isinstance(size, int) else size
		W = np.empty(shape=shape, dtype=np.float64, order="F")
		fill(rng, out=W)
		return W

	return _isotropic if size is None else _isotropic(size)


class ProbeStream:
	"""Device-resident view of the probe sequence `isotropic(pdf=..., seed=...)` would produce column by column."""

	def __init__(self, n: int, pdf: str = "rademacher", seed: Union[int, np.random.Generator, None] = None, dtype=np.float64):
		assert pdf in _ISO_DISTRIBUTIONS.keys(), f"Invalid distribution '{pdf}' supplied."
		self.n = int(n)
		self.pdf = _ISO_DISTRIBUTIONS[pdf]
		self.dtype = np.dtype(dtype)
		self.rng = np.random.default_rng(seed)
		bg = self.rng.bit_generator
		self.on_device = self.pdf == "rademacher" and isinstance(bg, np.random.PCG64)
		if self.on_device:
			st = bg.state["state"]
			self._state, self._inc = int(st["state"]), int(st["inc"])
		self.position = 0  # probes handed out so far

	def block(self, k: int, first: Optional[int] = None):
		"""Device probe block (n, k) for probes [first, first+k) (default: the next k); rows = vector entries."""
		from . import engine

		if self.on_device:
			first = self.position if first is None else int(first)
			X = engine.rademacher_block(self._state, self._inc, first, k, self.n, self.dtype)
			self.position = max(self.position, first + k)
			return X
		if first is not None and first > self.position:
			self.skip(first - self.position)  # another rank's probes: drawn and dropped to stay on the one stream
		assert first is None or first == self.position, "host-generated probe streams are sequential"
		W = np.empty((self.n, k), dtype=np.float64, order="F")
		_ISO_FUNCS[self.pdf](self.rng, W)
		self.position += k
		return engine.block_from_columns(W, self.dtype)

	def skip(self, count: int) -> None:
		"""Jumps over `count` probes (rank sharding)."""
		if self.on_device:
			self.position += int(count)
		else:
			for _ in range(int(count)):
				_ISO_FUNCS[self.pdf](self.rng, np.empty((self.n, 1), order="F"))
			self.position += int(count)

	def finish(self, consumed: int) -> None:
		"""Leaves the user's Generator where the reference would: `consumed` probes of n raw draws each."""
		if self.on_device:
			self.rng.bit_generator.advance(int(consumed) * self.n)


class Isotropic:
	"""Thread-pool filler of an (N, k) F-ordered array, one spawned sub-stream per thread (random.py:100-142)."""

	def __init__(self, size: tuple, pdf: str = "signs", seed: Union[int, np.random.SeedSequence, np.random.Generator, None] = None,
	             threads: Optional[int] = None):  # fmt: skip
		assert pdf in _ISO_DISTRIBUTIONS.keys(), f"Invalid distribution '{pdf}' supplied."
		self.pdf = _ISO_DISTRIBUTIONS[pdf]
		self.iso = _ISO_FUNCS[self.pdf]
		self.threads = multiprocessing.cpu_count() if threads is None else int(threads)
		rng = np.random.default_rng(seed)
		assert isinstance(rng, np.random.Generator) and hasattr(rng, "spawn")
		self._random_generators = [rng] if self.threads == 1 else rng.spawn(self.threads)
		self.shape = size
		self.executor = concurrent.futures.ThreadPoolExecutor(self.threads)
		self.values = np.zeros(size, order="F")
		self.step = np.ceil(self.shape[1] / self.threads).astype(np.int_)

	def fill(self):
		jobs = []
		for i, rng in enumerate(self._random_generators):
			lo, hi = i * self.step, (i + 1) * self.step
			jobs.append(self.executor.submit(lambda g, a, b: self.iso(g, out=self.values[:, a:b]), rng, lo, hi))
		concurrent.futures.wait(jobs)

	def __del__(self):
		self.executor.shutdown(False)


def symmetric(n: int, dist: str = "normal", pd: bool = False, ew: Optional[np.ndarray] = None,
              seed: Union[int, np.random.Generator, None] = None) -> np.ndarray:  # fmt: skip
	"""Random symmetric test matrix with prescribed (or random) eigenvalues `ew` (random.py:145-180)."""
	rng = np.random.default_rng() if seed is None else np.random.default_rng(seed)
	n_off = n * (n - 1) // 2
	if dist == "uniform":
		A = sp.spatial.distance.squareform(rng.uniform(size=n_off))
	elif dist == "normal":
		A = sp.spatial.distance.squareform(rng.normal(size=n_off))
	else:
		raise ValueError(f"Invalid distribution {dist} supplied")
	A[np.diag_indices(n)] = rng.random(n)
	Q, _ = np.linalg.qr(A)
	if ew is None:
		ew = rng.uniform(size=n, low=0.0 if pd else -1.0, high=1.0)
	B = Q @ np.diag(np.atleast_1d(ew)) @ Q.T
	return (B + B.T) / 2


def haar(n: int, ew: Optional[np.ndarray] = None, seed: Union[int, np.random.Generator, None] = None) -> np.ndarray:
	"""Random matrix with eigenvalues `ew` and Haar-distributed eigenvectors (random.py:183-201)."""
	rng = np.random.default_rng() if seed is None else np.random.default_rng(seed)
	group = sp.stats.ortho_group(n, seed=rng)
	ew = rng.uniform(size=n, low=-1.0, high=1.0) if ew is None else np.atleast_1d(ew)
	assert len(ew) == n, "Number of eigenvalues must be <= `n`"
	U = group.rvs()
	return U @ np.diag(ew) @ U.T
