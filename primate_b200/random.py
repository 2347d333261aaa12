"""Isotropic probe vectors (API mirror of src/primate/random.py:11-201).

`isotropic` keeps the reference's host contract (an F-ordered float64 (N, k) array filled in memory order from one
NumPy PCG64 stream, random.py:22-29,74-78).  The estimators do not call it on the hot path: they use
`ProbeStream`, which produces the SAME probes directly in GPU memory -- Rademacher probes by jump-ahead into the
PCG64 stream on the device (csrc/slq_probes.cu; probe i = stream elements [i*N, (i+1)*N), sign = top bit of the raw
output), normal/sphere probes by NumPy's own ziggurat restated on the device (csrc/slq_normals.cu: the data-dependent
number of raw draws per variate is resolved by a self-synchronising parallel scan).  Either way the
user's Generator ends up advanced exactly as the reference would leave it.
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


ZIG_R = 3.6541528853610087963519472518  # NumPy's ziggurat_nor_r / ziggurat_nor_inv_r (ziggurat_constants.h)
ZIG_INV_R = 0.27366123732975827203338247596
_SEEK_CHUNK = 1 << 27  # raw draws scanned per seek launch
_PLAN_CACHE: dict = {}


def _pairwise_plan(n: int):
	"""NumPy's pairwise summation of n contiguous doubles (numpy/_core/src/umath/loops_utils.h.src: pairwise_sum) as a
	list of leaf blocks (start, length <= 128) plus, per leaf, the number of pair-additions to perform after it on a
	stack machine -- what `np.sum(W**2, axis=0)` does per column of an F-ordered array (random.py:38)."""
	if n in _PLAN_CACHE:
		return _PLAN_CACHE[n]
	starts, lens, red = [], [], []
	stack = [(0, n, False)]  # (start, length, right-child-done marker)
	while stack:
		start, length, marker = stack.pop()
		if marker:
			red[-1] += 1
			continue
		if length <= 128:
			starts.append(start)
			lens.append(length)
			red.append(0)
			continue
		n2 = length // 2
		n2 -= n2 % 8
		stack.append((0, 0, True))  # after both children: one addition
		stack.append((start + n2, length - n2, False))
		stack.append((start, n2, False))
	plan = (np.asarray(starts, dtype=np.int64), np.asarray(lens, dtype=np.int32), np.asarray(red, dtype=np.int32))
	_PLAN_CACHE[n] = plan
	return plan


class ProbeStream:
	"""Device-resident view of the probe sequence `isotropic(pdf=..., seed=...)` would produce column by column.

	Rademacher probes are a pure function of (seed, probe index): random access by LCG jump-ahead.  Normal / sphere
	probes consume a data-dependent number of raw draws each (NumPy's ziggurat), so the stream position of probe i is
	found by a device scan (libslq: slq_normals_scan) from the nearest known position; the positions of the block
	boundaries met so far are remembered, so sequential use never scans twice."""

	def __init__(self, n: int, pdf: str = "rademacher", seed: Union[int, np.random.Generator, None] = None, dtype=np.float64):
		assert pdf in _ISO_DISTRIBUTIONS.keys(), f"Invalid distribution '{pdf}' supplied."
		self.n = int(n)
		self.pdf = _ISO_DISTRIBUTIONS[pdf]
		self.dtype = np.dtype(dtype)
		self.rng = np.random.default_rng(seed)
		bg = self.rng.bit_generator
		self.on_device = isinstance(bg, np.random.PCG64)
		if self.on_device:
			st = bg.state["state"]
			self._state, self._inc = int(st["state"]), int(st["inc"])
		self.position = 0  # probes handed out so far
		self._raw_at = {0: 0}  # normal / sphere: probe index -> raw draws consumed before it
		self._plan_dev = None

	# ---- normal / sphere on the device ------------------------------------------------------------------------------
	def _scan_ws(self, nraw: int):
		from . import _ffi, engine

		torch = engine._torch()
		need = _ffi.lib().slq_normals_workspace_bytes(int(nraw))
		ws = getattr(self, "_ws", None)
		if ws is None or ws.numel() < need:
			ws = self._ws = torch.empty(need, dtype=torch.uint8, device="cuda")
		return ws

	def _raw_offset(self, probe: int) -> int:
		"""Raw draws consumed by probes [0, probe) of the normal stream (exact)."""
		from . import _ffi, engine

		probe = int(probe)
		if probe in self._raw_at:
			return self._raw_at[probe]
		torch = engine._torch()
		L = _ffi.lib()
		start = max(p for p in self._raw_at if p <= probe)
		base, remaining = self._raw_at[start], (probe - start) * self.n
		total = torch.zeros(1, dtype=torch.int64, device="cuda")
		raw_out = torch.zeros(1, dtype=torch.int64, device="cuda")
		st, inc = _ffi.u128_pair(self._state), _ffi.u128_pair(self._inc)
		while remaining > 0:
			nraw = int(min(_SEEK_CHUNK, remaining * 1.03 + 8192))
			ws = self._scan_ws(nraw)
			# first launch: how many samples start in the chunk; second: locate the wanted one (or the chunk's last)
			_ffi.check(L.slq_normals_scan(st, inc, base, nraw, -1, total.data_ptr(), None, ws.data_ptr(), ws.numel(), engine._stream_ptr()))
			have = int(total.item())
			want = remaining if remaining < have else have - 1
			_ffi.check(L.slq_normals_scan(st, inc, base, nraw, want, total.data_ptr(), raw_out.data_ptr(), ws.data_ptr(), ws.numel(),
			                              engine._stream_ptr()))  # fmt: skip
			base += int(raw_out.item())
			remaining -= want
		self._raw_at[probe] = base
		return base

	def _normal_block(self, k: int, first: int):
		"""float64 device block (n, k) of standard normals for probes [first, first + k), bit-exact with NumPy."""
		import math

		from . import _ffi, engine

		torch = engine._torch()
		L = _ffi.lib()
		n = self.n
		base = self._raw_offset(first)
		X = torch.empty((n, k), dtype=torch.float64, device="cuda")
		end_raw = torch.zeros(1, dtype=torch.int64, device="cuda")
		tail_count = torch.zeros(1, dtype=torch.int64, device="cuda")
		cap = int(n * k * 1e-3) + 4096  # expected tail samples: 2.6e-4 n k
		tail_ord = torch.empty(cap, dtype=torch.int64, device="cuda")
		tail_raw = torch.empty(cap, dtype=torch.int64, device="cuda")
		st, inc = _ffi.u128_pair(self._state), _ffi.u128_pair(self._inc)
		nraw = int(n * k * 1.03) + 8192
		while True:
			ws = self._scan_ws(nraw)
			_ffi.check(L.slq_probes_normal(st, inc, base, nraw, int(k), n, X.data_ptr(), end_raw.data_ptr(), tail_count.data_ptr(), cap,
			                               tail_ord.data_ptr(), tail_raw.data_ptr(), ws.data_ptr(), ws.numel(), engine._stream_ptr()))  # fmt: skip
			used = int(end_raw.item())
			if used >= 0:  # UINT64_MAX (read as -1): the scan was too short for n*k samples
				break
			nraw = int(nraw * 1.05) + 65536
		self._raw_at[first + k] = base + used
		# tail samples with the host's libm, exactly as NumPy computes them (distributions.c: random_standard_normal)
		nt = int(tail_count.item())
		assert nt <= cap, "tail list overflow"
		if nt:
			ords = tail_ord[:nt].cpu().numpy()
			raws = tail_raw[:nt].cpu().numpy().view(np.uint64)
			vals = np.empty(nt)
			for i in range(nt):
				xx = -ZIG_INV_R * math.log1p(-(float(int(raws[i]) >> 11) * (1.0 / 9007199254740992.0)))
				vals[i] = -(ZIG_R + xx) if (int(ords[i]) & 1) else ZIG_R + xx
			o = ords >> 1
			flat = torch.from_numpy((o % n) * k + o // n).cuda()
			X.view(-1)[flat] = torch.from_numpy(vals).cuda()
		return X

	def _sphere_scale(self, X) -> None:
		from . import _ffi, engine

		torch = engine._torch()
		n, k = int(X.shape[0]), int(X.shape[1])
		if self._plan_dev is None:
			starts, lens, red = _pairwise_plan(n)
			self._plan_dev = (torch.from_numpy(starts).cuda(), torch.from_numpy(lens).cuda(), torch.from_numpy(red).cuda())
		starts, lens, red = self._plan_dev
		S = torch.empty(starts.numel() * k, dtype=torch.float64, device="cuda")
		norm2 = torch.empty(k, dtype=torch.float64, device="cuda")
		_ffi.check(_ffi.lib().slq_probes_sphere_scale(X.data_ptr(), n, k, starts.data_ptr(), lens.data_ptr(), red.data_ptr(),
		                                              int(starts.numel()), S.data_ptr(), norm2.data_ptr(), engine._stream_ptr()))  # fmt: skip

	def block(self, k: int, first: Optional[int] = None):
		"""Device probe block (n, k) for probes [first, first+k) (default: the next k); rows = vector entries."""
		from . import engine

		if self.on_device:
			first = self.position if first is None else int(first)
			if self.pdf == "rademacher":
				X = engine.rademacher_block(self._state, self._inc, first, k, self.n, self.dtype)
			else:
				X = self._normal_block(int(k), first)
				if self.pdf == "sphere":
					self._sphere_scale(X)
				if self.dtype != np.float64:
					X = X.to(engine._tdtype(self.dtype))  # the reference casts after generating float64 (trace.py:108,115)
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
		"""Leaves the user's Generator where the reference would: advanced by exactly the raw draws the first `consumed`
		probes use (n each for Rademacher; the scanned, data-dependent count for normal / sphere probes).  A generator
		that is not PCG64 was drawn from directly on the host and is already where the probes handed out left it."""
		if not self.on_device:
			return
		raws = int(consumed) * self.n if self.pdf == "rademacher" else self._raw_offset(int(consumed))
		self.rng.bit_generator.advance(raws)


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
