"""Device-side engine: operator handles and block calls into libslq.  torch tensors are device buffers only.

Everything here runs on the CUDA device or raises; nothing falls back to NumPy.
"""
from __future__ import annotations

import ctypes
import os
from typing import Any, Optional

import numpy as np

from . import _ffi

MAX_BLOCK = 64


def _torch():
	import torch

	if not torch.cuda.is_available():
		raise _ffi.SlqError("primate_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
	return torch


def _stream_ptr() -> int:
	return _torch().cuda.current_stream().cuda_stream


def _tdtype(np_dtype):
	torch = _torch()
	return torch.float64 if np.dtype(np_dtype) == np.float64 else torch.float32


def _code(np_dtype) -> int:
	dt = np.dtype(np_dtype)
	if dt == np.float64:
		return _ffi.SLQ_F64
	if dt == np.float32:
		return _ffi.SLQ_F32
	raise TypeError("Only 32- or 64-bit floats are supported.")  # mirrors the pybind11 overload failure


REORDER_GRID = 148  # CTAs of the persistent step-A grid (SMs of a B200)
REORDER_TILE = 120  # rows per cluster: the row tile of the step-A kernel at k = 32 fp64 (slq_stepa_tma.cuh)
REORDER_MIN_ROWS = 1 << 16  # below this every probe block fits in L2/L1 anyway


def _want_reorder(reorder, n: int, nnz: int) -> bool:
	"""`reorder`: "auto" (large, sparse operators), True/False, or an int cluster size.  SLQ_REORDER=0/1 overrides "auto"."""
	import os

	if reorder is None or reorder is False:
		return False
	if reorder == "auto":
		env = os.environ.get("SLQ_REORDER")
		if env is not None:
			return env not in ("0", "false", "off", "")
		return n >= REORDER_MIN_ROWS and nnz <= 64 * n
	return bool(reorder)


_ORDER_CACHE: dict = {}  # (n, nnz, tile, structure fingerprint) -> locality permutation (host int32 array); at most 4 entries


def _structure_key(indptr: np.ndarray, indices: np.ndarray, tile: int) -> tuple:
	"""Cheap fingerprint of a CSR sparsity structure: sizes + a hash of <= 64k evenly spaced entries of each index array.
	Used ONLY to re-use a locality ordering: any permutation is a valid one (P A P^T has the spectrum of A and the
	estimators are invariant under it), so a collision could cost locality, never correctness."""
	import hashlib

	h = hashlib.blake2b(digest_size=16)
	for a in (indptr, indices):
		step = max(1, len(a) // 65536)
		h.update(np.ascontiguousarray(a[::step]).tobytes())
	return (len(indptr) - 1, len(indices), int(tile), h.hexdigest())


def _cluster_order(L, n: int, ip: np.ndarray, ix: np.ndarray, tile: int) -> np.ndarray:
	"""slq_csr_cluster_order with the result remembered per sparsity structure (a MatrixFunction rebuilt on the same matrix
	-- another f, another degree, the next step of a solver loop -- does not repeat the breadth-first clustering)."""
	key = _structure_key(ip, ix, tile)
	perm = _ORDER_CACHE.get(key)
	if perm is None:
		perm = np.empty(n, dtype=np.int32)
		_ffi.check(L.slq_csr_cluster_order(n, ip.ctypes.data, ix.ctypes.data, tile, perm.ctypes.data))
		if len(_ORDER_CACHE) >= 4:
			_ORDER_CACHE.pop(next(iter(_ORDER_CACHE)))
		_ORDER_CACHE[key] = perm
	return perm


FUSED_MAX_OFFTILE = 1.75  # off-tile non-zeros per row up to which the one-kernel Lanczos step pays (slq_op_set_fused_step)


def _offtile_per_row(indptr: np.ndarray, indices: np.ndarray, tile: int = REORDER_TILE, sample: int = 512) -> float:
	"""Non-zeros per row whose column lies outside the row's tile of `tile` consecutive rows, estimated on `sample`
	evenly spaced tiles (host CSR arrays; about a millisecond whatever nnz is)."""
	n = len(indptr) - 1
	nt = -(-n // tile)
	if nt == 0 or len(indices) == 0:
		return 0.0
	ts = np.unique(np.linspace(0, nt - 1, min(nt, sample)).astype(np.int64))
	r0 = ts * tile
	r1 = np.minimum(r0 + tile, n)
	lo, hi = indptr[r0].astype(np.int64), indptr[r1].astype(np.int64)
	cnt = hi - lo
	total = int(cnt.sum())
	if total == 0:
		return 0.0
	start = np.cumsum(cnt) - cnt
	idx = np.repeat(lo - start, cnt) + np.arange(total, dtype=np.int64)
	c = indices[idx]
	off = int(np.count_nonzero((c < np.repeat(r0, cnt)) | (c >= np.repeat(r1, cnt))))
	return off / max(int((r1 - r0).sum()), 1)


def _offtile_per_row_device(rowptr, colidx, n: int, tile: int = REORDER_TILE, sample: int = 512) -> float:
	"""`_offtile_per_row` for CSR arrays that live on the GPU (index bookkeeping with torch on the sampled tiles only;
	one scalar is read back)."""
	torch = _torch()
	nt = -(-n // tile)
	if nt == 0 or colidx.numel() == 0:
		return 0.0
	dev = rowptr.device
	ts = torch.unique(torch.linspace(0, nt - 1, min(nt, sample), device=dev).long())
	r0 = ts * tile
	r1 = torch.clamp(r0 + tile, max=n)
	rp = rowptr.long()
	lo, hi = rp[r0], rp[r1]
	cnt = hi - lo
	total = int(cnt.sum().item())
	if total == 0:
		return 0.0
	start = torch.cumsum(cnt, 0) - cnt
	idx = torch.repeat_interleave(lo - start, cnt) + torch.arange(total, device=dev)
	c = colidx[idx].long()
	a, b = torch.repeat_interleave(r0, cnt), torch.repeat_interleave(r1, cnt)
	off = int(((c < a) | (c >= b)).sum().item())
	return off / max(int((r1 - r0).sum().item()), 1)


def _permute_csr_device(rowptr, colidx, vals, perm, n: int):
	"""P A P^T for CSR arrays on the GPU (perm[new] = old), rows sorted by their new column indices: the arrays
	slq_csr_permute produces on the host, bit for bit, without the host pass over nnz (index bookkeeping with torch:
	one radix sort of row * n + column keys).  Returns (rowptr int32, colidx int32, vals)."""
	import torch

	dev = rowptr.device
	rp = rowptr.long()
	deg = rp[1:] - rp[:-1]
	p = perm.long()
	inv = torch.empty(n, dtype=torch.int64, device=dev)
	inv[p] = torch.arange(n, device=dev)
	oldrow = torch.repeat_interleave(torch.arange(n, device=dev), deg)
	key = inv[oldrow] * n + inv[colidx.long()]
	del oldrow
	key, order = torch.sort(key)
	ncol = (key % n).int()
	del key
	nval = vals[order]
	nrp = torch.zeros(n + 1, dtype=torch.int64, device=dev)
	nrp[1:] = torch.cumsum(deg[p], 0)
	return nrp.int(), ncol, nval


def _balance_tiles(weights: np.ndarray, ncta: int) -> np.ndarray:
	"""Tile placement for a persistent grid: returns `place` with place[j] = the tile to put at position j, where
	position j is processed by CTA j % ncta.  Longest-processing-time greedy with the per-CTA tile counts fixed (heaviest
	tile to the least loaded CTA that still has a free position); each CTA runs its tiles heaviest first.  The last
	tile keeps its position (it may be short of rows)."""
	import heapq

	nt = len(weights)
	if nt < 2 * ncta:
		return np.arange(nt)
	cap = [len(range(c, nt - 1, ncta)) for c in range(ncta)]  # free positions per CTA, the last position excluded
	load = [0] * ncta
	load[(nt - 1) % ncta] = int(weights[nt - 1])
	heap = [(load[c], c) for c in range(ncta) if cap[c] > 0]
	heapq.heapify(heap)
	got = [[] for _ in range(ncta)]
	for t in np.argsort(-weights[: nt - 1], kind="stable"):
		ld, c = heapq.heappop(heap)
		got[c].append(int(t))
		cap[c] -= 1
		if cap[c] > 0:
			heapq.heappush(heap, (ld + int(weights[t]), c))
	place = np.empty(nt, dtype=np.int64)
	place[nt - 1] = nt - 1
	for c in range(ncta):
		place[c : nt - 1 : ncta] = got[c]
	return place


def _balance_rows_device(rowptr, colidx, vals, n: int, tile: int):
	"""P A P^T with the rows dealt round-robin to the row tiles by decreasing length, so that every tile (hence every
	persistent CTA of the step-A kernel) carries about the same number of non-zeros.  Device tensors in and out
	(index bookkeeping with torch; the SpMM itself is libslq's).  Returns (rowptr, colidx, vals, perm[new] = old)."""
	torch = _torch()
	dev = rowptr.device
	rp = rowptr.long()
	deg = rp[1:] - rp[:-1]
	order = torch.argsort(deg, descending=True, stable=True)
	ntiles = -(-n // tile)
	pad = torch.full((ntiles * tile - n,), -1, dtype=torch.int64, device=dev)
	grid = torch.cat([order, pad]).reshape(tile, ntiles)  # rank r -> slot r // ntiles of tile r % ntiles
	# Tile t is processed by persistent CTA t % 148.  The tiles are equal but for their one hub row, and the top hubs
	# alone are ~10% of a CTA's non-zeros on an R-MAT graph: place the tiles so that every CTA gets the same total.
	weights = torch.where(grid >= 0, deg[grid.clamp(min=0)], torch.zeros_like(grid)).sum(dim=0)
	place = torch.from_numpy(_balance_tiles(weights.cpu().numpy(), REORDER_GRID)).to(dev)
	grid = grid[:, place].t().reshape(-1)
	perm = grid[grid >= 0]
	inv = torch.empty(n, dtype=torch.int64, device=dev)
	inv[perm] = torch.arange(n, device=dev)
	ndeg = deg[perm]
	nrp = torch.zeros(n + 1, dtype=torch.int64, device=dev)
	nrp[1:] = torch.cumsum(ndeg, 0)
	nnz = int(nrp[-1].item())
	src = torch.repeat_interleave(rp[:-1][perm] - nrp[:-1], ndeg) + torch.arange(nnz, device=dev)
	ncol = inv[colidx.long()[src]].int()
	nval = vals[src]
	return nrp.int(), ncol, nval, perm.int()


def _device_csr_from(A, dtype, dev):
	"""COO or BSR SciPy matrix -> (rowptr, colidx, vals) int32/int32/float device tensors with sorted, de-duplicated
	(summed, as SciPy's tocsr() does) column indices; the conversion runs on the GPU (index bookkeeping with torch: a
	radix sort on row*n + col keys and a segmented sum), only the raw format arrays cross PCIe.  Returns the bytes uploaded."""
	torch = _torch()
	n = int(A.shape[0])
	tdt = _tdtype(dtype)
	if A.format == "coo":
		r = torch.from_numpy(np.ascontiguousarray(A.row, dtype=np.int64)).to(dev)
		c = torch.from_numpy(np.ascontiguousarray(A.col, dtype=np.int64)).to(dev)
		v = torch.from_numpy(np.ascontiguousarray(A.data, dtype=np.dtype(dtype))).to(dev)
		up = A.row.nbytes + A.col.nbytes + A.data.nbytes
	else:  # bsr: block (bi, bj) at position p holds A[bi R + a, bj C + b] = data[p, a, b]
		R, Cb = A.blocksize
		ip = torch.from_numpy(np.ascontiguousarray(A.indptr, dtype=np.int64)).to(dev)
		bj = torch.from_numpy(np.ascontiguousarray(A.indices, dtype=np.int64)).to(dev)
		blocks = torch.from_numpy(np.ascontiguousarray(A.data, dtype=np.dtype(dtype))).to(dev)
		up = A.indptr.nbytes + A.indices.nbytes + A.data.nbytes
		bi = torch.repeat_interleave(torch.arange(ip.numel() - 1, device=dev), ip[1:] - ip[:-1])
		a = torch.arange(R, device=dev).view(1, R, 1)
		b = torch.arange(Cb, device=dev).view(1, 1, Cb)
		r = (bi.view(-1, 1, 1) * R + a).expand(-1, R, Cb).reshape(-1)
		c = (bj.view(-1, 1, 1) * Cb + b).expand(-1, R, Cb).reshape(-1)
		v = blocks.reshape(-1)
	key, order = torch.sort(r * n + c)
	v = v[order].to(tdt)
	uniq, inverse = torch.unique_consecutive(key, return_inverse=True)
	vals = torch.zeros(uniq.numel(), dtype=tdt, device=dev).index_add_(0, inverse, v)
	rows = uniq // n
	rowptr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
	rowptr[1:] = torch.cumsum(torch.bincount(rows, minlength=n), 0)
	assert uniq.numel() < 2**31 - 1, "nnz must fit int32"
	return rowptr.int(), (uniq % n).int(), vals, int(up)


class _DevView:
	"""Exposes a raw device pointer through __cuda_array_interface__ so torch can view it without a copy."""

	def __init__(self, ptr: int, shape: tuple, typestr: str):
		self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None}


def block_sizes(count: int, kmax: int = 32) -> list:
	"""Splits `count` probes into power-of-two blocks (largest first), e.g. 200 -> 6x32 + 8."""
	out, rest = [], int(count)
	while rest > 0:
		k = kmax
		while k > rest:
			k >>= 1
		out.append(k)
		rest -= k
	return out


class DeviceOperator:
	"""A symmetric operator resident on the GPU (slq_op handle).

	Accepts what the reference's FFI accepts (src/primate/_lanczos.cpp:104-111): a dense ndarray, a SciPy sparse
	matrix/array (any format; converted to CSR once, not per call), or any object with `.matvec` and `.shape`
	(slow host-callback path, include/pylinop.h:17-40).  The matrix is uploaded ONCE per operator.

	`fused` (sparse input): "auto" lets the one-kernel Lanczos step (DESIGN 3 "fused step") run where it pays -- at most
	1.75 non-zeros per row outside their 120-row tile, measured on the final row order -- True / False force it.

	`tiled` (dense fp64 input): keep a band-major copy of the matrix for the tensor-core product kernel ("auto": when it
	fits comfortably; False: row-major only; results are bit-identical either way).

	`reorder` (sparse input): "auto" picks a symmetric row/column order for large operators -- breadth-first locality
	clusters of one SpMM row tile for meshes and the like, or, when the row lengths are heavily skewed (graph hubs),
	rows dealt to the tiles by decreasing length so that every tile carries the same number of non-zeros
	("balance").  True / an int cluster size / "balance" force one of them, False keeps the caller's order.  The
	order is internal: every vector that crosses the API is in the caller's row order.
	"""

	def __init__(self, A: Any, dtype=None, reorder="auto", tiled="auto", fused="auto"):
		import scipy.sparse as sp

		torch = _torch()
		L = _ffi.lib()
		self.shape = tuple(A.shape)
		assert len(self.shape) == 2 and self.shape[0] == self.shape[1], "This function only works with square, symmetric matrices!"
		self.n = int(self.shape[0])
		if dtype is None:
			dtype = A.dtype if hasattr(A, "dtype") else np.float64
		self.dtype = np.dtype(dtype)
		self.code = _code(self.dtype)
		self._keep = []
		self._handle = ctypes.c_void_p()
		self.h2d_bytes = 0
		self.tiled = False
		self.fused_step, self.offtile_per_row = True, None  # (the library's default wherever this class does not decide)
		dev = torch.device("cuda", torch.cuda.current_device())
		if sp.issparse(A) and A.format in ("coo", "bsr") and not (isinstance(reorder, int) and not isinstance(reorder, bool) and reorder > 1) \
				and reorder is not True and not _want_reorder(reorder, self.n, int(A.nnz)):
			# COO / BSR: the raw arrays are uploaded as they are and turned into sorted CSR ON THE DEVICE (no host tocsr())
			rowptr, colidx, vals, up = _device_csr_from(A, self.dtype, dev)
			self.h2d_bytes = up
			self.perm = None
			lens = rowptr[1:] - rowptr[:-1]
			if isinstance(reorder, str) and reorder == "balance":
				rowptr, colidx, vals, self.perm = _balance_rows_device(rowptr, colidx, vals, self.n, REORDER_TILE)
			self._keep = [rowptr, colidx, vals]
			self.kind, self.nnz = "csr", int(colidx.numel())
			_ffi.check(L.slq_csr_create_ex(self.n, self.nnz, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr(), self.code,
			                               int(lens.max().item()) if self.n else 0, None, ctypes.byref(self._handle)))  # fmt: skip
			if isinstance(fused, str) and fused == "auto":
				self.offtile_per_row = float("inf") if self.perm is not None else _offtile_per_row_device(rowptr, colidx, self.n)
				self.fused_step = self.offtile_per_row <= FUSED_MAX_OFFTILE
			else:
				self.fused_step = bool(fused)
			_ffi.check(L.slq_op_set_fused_step(self._handle, int(self.fused_step)))
		elif sp.issparse(A):
			if A.format == "csc":
				# a symmetric matrix in CSC form IS its CSR form (A^T = A): no conversion, exactly the arrays the reference's
				# pybind11 caster hands to Eigen (_lanczos.cpp:89-93)
				C = sp.csr_matrix((A.data, A.indices, A.indptr), shape=A.shape, copy=False)
				if not A.has_sorted_indices:
					C = C.copy()
			else:
				C = A.tocsr()
			C.sort_indices()
			assert C.nnz < 2**31 - 1, "nnz must fit int32"
			ip = np.ascontiguousarray(C.indptr, dtype=np.int32)
			ix = np.ascontiguousarray(C.indices, dtype=np.int32)
			va = np.ascontiguousarray(C.data, dtype=self.dtype)
			self.perm = None
			# power-law row lengths (hubs): no locality to find; deal the rows to the tiles by length instead (on the device)
			skewed = self.n >= REORDER_MIN_ROWS and int(np.diff(ip).max(initial=0)) > max(1024, 64 * int(C.nnz) // max(self.n, 1))
			balance = (isinstance(reorder, str) and reorder == "balance") or (isinstance(reorder, str) and reorder == "auto" and skewed and os.environ.get("SLQ_REORDER") is None)
			if not balance and _want_reorder(reorder, self.n, int(C.nnz)):
				# locality ordering (slq_csr_cluster_order): rows grouped into breadth-first clusters of one SpMM row tile
				tile = int(reorder) if (isinstance(reorder, int) and not isinstance(reorder, bool) and reorder > 1) else int(os.environ.get("SLQ_REORDER_TILE", REORDER_TILE))
				# (the order comes from the host -- breadth-first clustering, remembered per structure -- but it is APPLIED on
				# the device: the caller's arrays are uploaded as they are, no host pass over nnz)
				self.perm = torch.from_numpy(_cluster_order(L, self.n, ip, ix, tile)).to(dev)
			rowptr = torch.from_numpy(ip).to(dev, non_blocking=False)
			colidx = torch.from_numpy(ix).to(dev)
			vals = torch.from_numpy(va).to(dev)
			self.h2d_bytes = rowptr.numel() * 4 + colidx.numel() * 4 + vals.numel() * vals.element_size()
			if self.perm is not None:
				rowptr, colidx, vals = _permute_csr_device(rowptr, colidx, vals, self.perm, self.n)
			if balance:
				rowptr, colidx, vals, self.perm = _balance_rows_device(rowptr, colidx, vals, self.n, REORDER_TILE)
			self._keep = [rowptr, colidx, vals]
			if self.perm is not None and not balance:
				self.h2d_bytes += self.perm.numel() * 4
			self.kind, self.nnz = "csr", int(C.nnz)
			# the longest row is known on the host (a symmetric permutation keeps the row lengths): no device work, no sync
			_ffi.check(L.slq_csr_create_ex(self.n, self.nnz, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr(), self.code,
			                               int(np.diff(C.indptr).max(initial=0)), None, ctypes.byref(self._handle)))  # fmt: skip
			# one-kernel Lanczos step (orth = 0): it reads two rows per gathered non-zero, so it is kept for operators whose
			# non-zeros mostly stay inside a row tile (ip / ix are the arrays in the operator's final row order)
			self.offtile_per_row = float("inf") if balance else (_offtile_per_row(ip, ix) if self.perm is None else _offtile_per_row_device(rowptr, colidx, self.n))
			self.fused_step = (self.offtile_per_row <= FUSED_MAX_OFFTILE) if (isinstance(fused, str) and fused == "auto") else bool(fused)
			_ffi.check(L.slq_op_set_fused_step(self._handle, int(self.fused_step)))
		elif isinstance(A, np.ndarray) or (hasattr(A, "is_cuda") and hasattr(A, "data_ptr")):
			M = A if hasattr(A, "data_ptr") else torch.from_numpy(np.ascontiguousarray(A, dtype=self.dtype))
			M = M.to(dev, dtype=_tdtype(self.dtype))
			if not (M.stride(1) == 1 and M.stride(0) >= self.n and (M.stride(0) * M.element_size()) % 16 == 0):
				M = M.contiguous()  # (a row-major view with a row stride > n, e.g. columns [0, n) of a wider buffer, is used as it is)
			ld = int(M.stride(0))
			self.h2d_bytes = M.numel() * M.element_size()
			self.kind, self.nnz = "dense", self.n * self.n
			if self.n % 16 and self.n >= 16:
				# order not a multiple of 16: zero-pad once on the device so that the bulk-copy / tensor-core kernels apply
				npad = -(-self.n // 16) * 16
				P = torch.zeros((npad, npad), dtype=M.dtype, device=dev)
				P[: self.n, : self.n] = M
				self._keep = [P]
				_ffi.check(L.slq_dense_create_padded(self.n, npad, P.data_ptr(), npad, self.code, ctypes.byref(self._handle)))
			else:
				self._keep = [M]
				_ffi.check(L.slq_dense_create(self.n, M.data_ptr(), ld, self.code, ctypes.byref(self._handle)))
			# fp64: band-major copy for the tensor-core kernel (one contiguous bulk copy per stage instead of 16 strided lines;
			# the row-major original stays for the narrow-block kernels).  "auto": when the copy fits in a quarter of free HBM.
			need = ctypes.c_size_t(0)
			_ffi.check(L.slq_dense_tiled_bytes(self._handle, ctypes.byref(need)))
			if tiled and need.value and (tiled is True or need.value <= torch.cuda.mem_get_info()[0] // 4):
				tb = torch.empty(need.value, dtype=torch.uint8, device=dev)
				self._keep.append(tb)
				_ffi.check(L.slq_dense_set_tiled(self._handle, tb.data_ptr(), tb.numel(), _stream_ptr()))
				self.tiled = True
			if self.n <= 16384:  # few row tiles: let the dense kernels split the reduction dimension (16 slabs of n x 64)
				sc = torch.empty(16 * self.n * MAX_BLOCK * M.element_size(), dtype=torch.uint8, device=dev)
				self._keep.append(sc)
				_ffi.check(L.slq_op_set_scratch(self._handle, sc.data_ptr(), sc.numel()))
		else:
			if not hasattr(A, "matvec"):
				raise ValueError("Supplied object is missing 'matvec' attribute.")  # pylinop.h:24
			if not hasattr(A, "shape"):
				raise ValueError("Supplied object is missing 'shape' attribute.")  # pylinop.h:25
			self.kind, self.nnz = "callback", 0
			self._A = A
			self._cb = _ffi.MATMAT_FN(self._callback)
			self._cb_error = None
			_ffi.check(L.slq_callback_create(self.n, self.code, self._cb, None, ctypes.byref(self._handle)))
		self._ws = {}
		if not hasattr(self, "perm"):
			self.perm = None

	@classmethod
	def from_device_csr(cls, rowptr, colidx, vals, n: int, dtype=None, reorder=False, fused="auto"):
		"""Operator from CSR arrays that already live on the GPU (torch tensors: int32 rowptr (n+1), int32 colidx, float
		vals; column indices sorted per row recommended).  No host copy is made unless a locality ordering is asked for
		(`reorder=True` or a cluster size): the clustering runs on the host on a copy of the index arrays
		(slq_csr_cluster_order), the permutation is applied on the device.
		`reorder="balance"` is the ordering for power-law graphs (no locality to find, but hub rows that would pile up in a
		few row tiles): rows are dealt to the row tiles in order of decreasing length, entirely on the device."""
		torch = _torch()
		L = _ffi.lib()
		self = cls.__new__(cls)
		self.n, self.shape = int(n), (int(n), int(n))
		self.dtype = np.dtype(dtype) if dtype is not None else np.dtype(str(vals.dtype).replace("torch.", ""))
		self.code = _code(self.dtype)
		rowptr = rowptr.to(torch.int32).contiguous()
		colidx = colidx.to(torch.int32).contiguous()
		vals = vals.to(_tdtype(self.dtype)).contiguous()
		assert rowptr.is_cuda and colidx.is_cuda and vals.is_cuda and rowptr.numel() == self.n + 1
		self.perm, self.h2d_bytes = None, 0
		self.tiled, self.fused_step, self.offtile_per_row = False, True, None
		if isinstance(reorder, str) and reorder == "balance":
			rowptr, colidx, vals, self.perm = _balance_rows_device(rowptr, colidx, vals, self.n, REORDER_TILE)
		elif reorder:
			tile = int(reorder) if (isinstance(reorder, int) and not isinstance(reorder, bool) and reorder > 1) else REORDER_TILE
			# only the index structure visits the host (for the clustering); the permutation is applied on the device
			ip, ix = rowptr.cpu().numpy(), colidx.cpu().numpy()
			self.perm = torch.from_numpy(_cluster_order(L, self.n, ip, ix, tile)).to(rowptr.device)
			rowptr, colidx, vals = _permute_csr_device(rowptr, colidx, vals, self.perm, self.n)
		self._keep = [rowptr, colidx, vals]
		self.kind, self.nnz = "csr", int(colidx.numel())
		self._handle = ctypes.c_void_p()
		# longest row computed on the CURRENT stream (the arrays may still be in flight there) and read back: documented sync
		_ffi.check(L.slq_csr_create_ex(self.n, self.nnz, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr(), self.code,
		                               -1, _stream_ptr(), ctypes.byref(self._handle)))  # fmt: skip
		if isinstance(fused, str) and fused == "auto":
			balanced = isinstance(reorder, str) and reorder == "balance"
			self.offtile_per_row = float("inf") if balanced else _offtile_per_row_device(rowptr, colidx, self.n)
			self.fused_step = self.offtile_per_row <= FUSED_MAX_OFFTILE
		else:
			self.fused_step = bool(fused)
		_ffi.check(L.slq_op_set_fused_step(self._handle, int(self.fused_step)))
		self._ws = {}
		return self

	# ---- row order: device panels handed to / returned by this class are in the CALLER's row order ---------------------
	def _to_internal(self, X):
		"""Caller-order probe block -> the operator's (locality) row order.  Identity without a permutation."""
		if self.perm is None:
			return X
		torch = _torch()
		Xp = torch.empty_like(X)
		_ffi.check(_ffi.lib().slq_block_gather_rows(X.data_ptr(), self.perm.data_ptr(), self.n, int(X.shape[-1]), self.code,
		                                            Xp.data_ptr(), _stream_ptr()))  # fmt: skip
		return Xp

	def _from_internal(self, Yp):
		if self.perm is None:
			return Yp
		torch = _torch()
		Y = torch.empty_like(Yp)
		panels = Yp.reshape(-1, self.n, Yp.shape[-1])
		outs = Y.reshape(-1, self.n, Yp.shape[-1])
		for i in range(panels.shape[0]):
			_ffi.check(_ffi.lib().slq_block_scatter_rows(panels[i].data_ptr(), self.perm.data_ptr(), self.n, int(Yp.shape[-1]), self.code,
			                                             outs[i].data_ptr(), _stream_ptr()))  # fmt: skip
		return Y

	# host-callback operator: device panels out to the host, `A.matvec` per column, result back (pylinop.h:32-40)
	def _callback(self, ctx, x_ptr, y_ptr, n, k, dtype, stream) -> int:
		try:
			torch = _torch()
			ts = "<f8" if dtype == _ffi.SLQ_F64 else "<f4"
			X = torch.as_tensor(_DevView(x_ptr, (n, k), ts), device="cuda")
			Y = torch.as_tensor(_DevView(y_ptr, (n, k), ts), device="cuda")
			if hasattr(self._A, "matmat_device"):  # GPU-native user operator: torch panels in, torch panels out
				Y.copy_(self._A.matmat_device(X))
				return 0
			Xh = X.cpu().numpy()
			Yh = np.empty_like(Xh)
			for j in range(k):
				Yh[:, j] = np.asarray(self._A.matvec(np.ascontiguousarray(Xh[:, j])), dtype=Xh.dtype).ravel()[:n]
			Y.copy_(torch.from_numpy(Yh))
			return 0
		except Exception as exc:  # surfaced by the caller after the C call returns SLQ_ERR_CALLBACK
			self._cb_error = exc
			return 1

	def __del__(self):
		try:
			if getattr(self, "_handle", None) is not None and self._handle.value:
				_ffi.lib().slq_op_destroy(self._handle)
				self._handle = ctypes.c_void_p()
		except Exception:
			pass

	@property
	def handle(self):
		return self._handle

	def workspace(self, k: int, ncv: int):
		key = (k, ncv)
		if key not in self._ws:
			torch = _torch()
			nbytes = _ffi.lib().slq_lanczos_workspace_bytes(self.n, k, ncv, self.code)
			self._ws = {key: torch.empty(nbytes + 256, dtype=torch.uint8, device="cuda")}  # keep only the latest shape
		buf = self._ws[key]
		off = (-buf.data_ptr()) % 256
		return buf.data_ptr() + off, buf.numel() - off

	def _check(self, rc: int):
		if rc == _ffi.SLQ_ERR_CALLBACK and getattr(self, "_cb_error", None) is not None:
			exc, self._cb_error = self._cb_error, None
			raise exc
		_ffi.check(rc)

	def matvec(self, x: np.ndarray) -> np.ndarray:
		"""Host-array product (makes a DeviceOperator a valid `A` for MatrixFunction / hutch, linear_operator.h:25-29)."""
		x = np.asarray(x, dtype=self.dtype).reshape(self.n, -1)
		return block_to_columns(self.matmat(block_from_columns(x, self.dtype))).reshape(-1) if x.shape[1] == 1 else \
			block_to_columns(self.matmat(block_from_columns(x, self.dtype)))

	def matmat(self, X):
		"""Y = A X for a device probe block X (n, k) row-major (torch tensor)."""
		torch = _torch()
		X = self._to_internal(X)
		Y = torch.empty_like(X)
		self._check(_ffi.lib().slq_op_matmat(self._handle, X.data_ptr(), Y.data_ptr(), int(X.shape[1]), _stream_ptr()))
		return self._from_internal(Y)

	def lanczos(self, X, deg: int, rtol: float = 1e-8, orth: int = 0, ncv: int = 2, want_basis: bool = False, out: Optional[dict] = None,
	            internal_basis: bool = False):
		"""Blocked Lanczos on the device block X (n, k); X is clobbered.  Returns dict of device tensors.

		`out` may carry pre-allocated contiguous device tensors "alpha", "beta" (k, deg+1) and "xnorm2" (k,) -- e.g. row
		slices of one array holding the tridiagonals of many blocks -- in which case nothing is allocated here and the
		per-probe step counts are not returned."""
		torch = _torch()
		n, k = int(X.shape[0]), int(X.shape[1])
		assert n == self.n and X.is_contiguous() and X.dtype == _tdtype(self.dtype)
		if out is not None:
			alpha, beta, xnorm2, steps = out["alpha"], out["beta"], out["xnorm2"], out.get("steps")
			assert alpha.shape == (k, deg + 1) and beta.shape == (k, deg + 1) and alpha.is_contiguous() and beta.is_contiguous()
			assert alpha.dtype == X.dtype and xnorm2.dtype == torch.float64 and xnorm2.is_contiguous()
		else:
			alpha = torch.empty((k, deg + 1), dtype=X.dtype, device=X.device)
			beta = torch.empty((k, deg + 1), dtype=X.dtype, device=X.device)
			steps = torch.empty(k, dtype=torch.int32, device=X.device)
			xnorm2 = torch.empty(k, dtype=torch.float64, device=X.device)
		Q = torch.empty((ncv, n, k), dtype=X.dtype, device=X.device) if want_basis else None
		X = self._to_internal(X)  # (a copy when the operator carries a row permutation; the caller's X is then left intact)
		ws_ptr, ws_bytes = self.workspace(k, ncv)
		self._check(_ffi.lib().slq_lanczos(self._handle, X.data_ptr(), k, int(deg), float(rtol), int(orth), int(ncv), alpha.data_ptr(),
		                                   beta.data_ptr(), steps.data_ptr() if steps is not None else None, xnorm2.data_ptr(),
		                                   Q.data_ptr() if want_basis else None, ws_ptr, ws_bytes, _stream_ptr()))  # fmt: skip
		if want_basis and not internal_basis:  # internal_basis: Q stays in the operator's row order (see _from_internal)
			Q = self._from_internal(Q)
		return {"alpha": alpha, "beta": beta, "steps": steps, "xnorm2": xnorm2, "Q": Q}


	def _scratch(self, key, nbytes: int):
		"""256-byte aligned device scratch of at least `nbytes`, cached under `key` (only the latest shape is kept)."""
		if key not in self._ws:
			self._ws = {key: _torch().empty(nbytes + 256, dtype=_torch().uint8, device="cuda")}
		buf = self._ws[key]
		off = (-buf.data_ptr()) % 256
		return buf.data_ptr() + off, buf.numel() - off

	def funm_matmat(self, X, deg: int, rtol: float, fun_id: int, params):
		"""Y = f(A) X for a device probe block X (n, k) in the CALLER's row order, left intact: the basis-free two-pass
		Lanczos engine (libslq: slq_funm_matmat; orth = 0, built-in f).  Five panels of memory whatever `deg` is."""
		torch = _torch()
		L = _ffi.lib()
		n, k = int(X.shape[0]), int(X.shape[1])
		assert n == self.n and X.is_contiguous() and X.dtype == _tdtype(self.dtype)
		Xi = self._to_internal(X)
		Y = torch.empty_like(Xi)
		ws_ptr, ws_bytes = self._scratch(("funm", k, int(deg)), L.slq_funm_workspace_bytes(self.n, k, int(deg), self.code))
		self._check(L.slq_funm_matmat(self._handle, Xi.data_ptr(), k, int(deg), float(rtol), int(fun_id), params, Y.data_ptr(), ws_ptr,
		                              ws_bytes, _stream_ptr()))  # fmt: skip
		return self._from_internal(Y)

	def lanczos_replay(self, X, deg: int, run: dict, coeffs):
		"""Second pass of the two-pass engine with caller-provided Krylov coefficients `coeffs` (k, deg) float64 device
		tensor: `run` is what `lanczos(copy of X, deg, rtol, orth=0, ncv=2)` returned for the same start block."""
		torch = _torch()
		L = _ffi.lib()
		n, k = int(X.shape[0]), int(X.shape[1])
		Xi = self._to_internal(X)
		Y = torch.empty_like(Xi)
		ws_ptr, ws_bytes = self._scratch(("replay", k, int(deg)), L.slq_replay_workspace_bytes(self.n, k, int(deg), self.code))
		self._check(L.slq_lanczos_replay(self._handle, Xi.data_ptr(), k, int(deg), run["alpha"].data_ptr(), run["beta"].data_ptr(),
		                                 run["steps"].data_ptr(), run["xnorm2"].data_ptr(), coeffs.contiguous().data_ptr(), Y.data_ptr(),
		                                 ws_ptr, ws_bytes, _stream_ptr()))  # fmt: skip
		return self._from_internal(Y)


def fun_spec(fun, kwargs: Optional[dict] = None):
	"""(fun_id, params) of a built-in spectral function name (special.py:78-107), or None for a Python callable."""
	kw = dict(kwargs or {})
	if fun is None or fun == "identity":
		return 0, _ffi.fun_params()
	if not isinstance(fun, str):
		return None
	if fun == "exp":
		return 5, _ffi.fun_params(float(kw.get("t", 1.0)))
	if fun == "smoothstep":
		return 6, _ffi.fun_params(float(kw.get("a", 0.0)), float(kw.get("b", 1.0)))
	if fun == "numrank":
		return 7, _ffi.fun_params(float(kw.get("threshold", 0.000001)))
	if fun == "softsign":
		return 8, _ffi.fun_params(float(kw.get("q", 10)))
	if fun in _ffi.FUN_IDS:
		return _ffi.FUN_IDS[fun], _ffi.fun_params()
	raise ValueError(f"Unknown function: {fun}.")


def quadrature(alpha, beta, deg: int, fun_id: int = -1, params=None, scale=None, want_nodes: bool = True, want_y: bool = True):
	"""Batched Golub-Welsch on device tensors alpha/beta (k, ld).  Returns (theta, tau, y, info) device tensors."""
	torch = _torch()
	k, ld = int(alpha.shape[0]), int(alpha.shape[1])
	code = _ffi.SLQ_F64 if alpha.dtype == torch.float64 else _ffi.SLQ_F32
	theta = torch.empty((k, deg), dtype=torch.float64, device=alpha.device) if want_nodes else None
	tau = torch.empty((k, deg), dtype=torch.float64, device=alpha.device) if want_nodes else None
	y = torch.empty(k, dtype=torch.float64, device=alpha.device) if (want_y and fun_id >= 0) else None
	info = torch.empty(k, dtype=torch.int32, device=alpha.device)
	_ffi.check(_ffi.lib().slq_quadrature(alpha.data_ptr(), beta.data_ptr(), code, k, int(deg), ld, int(fun_id),
	                                     params if params is not None else _ffi.fun_params(),
	                                     scale.data_ptr() if scale is not None else None,
	                                     theta.data_ptr() if theta is not None else None, tau.data_ptr() if tau is not None else None,
	                                     y.data_ptr() if y is not None else None, info.data_ptr(), _stream_ptr()))  # fmt: skip
	return theta, tau, y, info


def warn_unconverged(info, where: str) -> None:
	"""Host-facing solvers (which synchronise anyway) report eigenvalues that ran out of QL sweeps; the reference's tqli
	tolerates the same silently (tqli.py:54-56), its LAPACK path would raise and fall back (tridiag.py:16-20)."""
	bad = int((info > 0).sum().item())
	if bad:
		import warnings

		warnings.warn(f"{where}: {bad} Jacobi matrices had eigenvalues that did not converge within the QL sweep limit", RuntimeWarning, stacklevel=3)


def funm_coeffs(alpha, beta, deg: int, fun_id: int, params=None):
	"""Krylov coefficients c_p = Y f(Theta) Y^T e_1 (k, deg) plus (theta, tau)."""
	torch = _torch()
	k, ld = int(alpha.shape[0]), int(alpha.shape[1])
	code = _ffi.SLQ_F64 if alpha.dtype == torch.float64 else _ffi.SLQ_F32
	kw = dict(dtype=torch.float64, device=alpha.device)
	coeffs, theta, tau = torch.empty((k, deg), **kw), torch.empty((k, deg), **kw), torch.empty((k, deg), **kw)
	scratch = torch.empty(k * deg * deg + k * deg, **kw)
	info = torch.empty(k, dtype=torch.int32, device=alpha.device)
	_ffi.check(_ffi.lib().slq_funm_coeffs(alpha.data_ptr(), beta.data_ptr(), code, k, int(deg), ld, int(fun_id),
	                                      params if params is not None else _ffi.fun_params(), coeffs.data_ptr(), theta.data_ptr(),
	                                      tau.data_ptr(), scratch.data_ptr(), info.data_ptr(), _stream_ptr()))  # fmt: skip
	return coeffs, theta, tau, scratch, info


def rademacher_block(state: int, inc: int, first_probe: int, k: int, n: int, dtype):
	"""Device Rademacher block (n, k) for probes first_probe .. first_probe+k-1 of the PCG64 stream (state, inc)."""
	torch = _torch()
	X = torch.empty((n, k), dtype=_tdtype(dtype), device="cuda")
	_ffi.check(_ffi.lib().slq_probes_rademacher(_ffi.u128_pair(state), _ffi.u128_pair(inc), int(first_probe), int(k), int(n),
	                                            _code(dtype), X.data_ptr(), _stream_ptr()))  # fmt: skip
	return X


def pcg64_raw(state: int, inc: int, offset: int, count: int):
	torch = _torch()
	out = torch.empty(count, dtype=torch.int64, device="cuda")
	_ffi.check(_ffi.lib().slq_pcg64_raw(_ffi.u128_pair(state), _ffi.u128_pair(inc), int(offset), int(count), out.data_ptr(), _stream_ptr()))
	return out.cpu().numpy().view(np.uint64)


def block_from_columns(W: np.ndarray, dtype):
	"""Uploads an (n, k) host panel (any order) and returns the device probe block (n, k) row-major."""
	torch = _torch()
	Wf = np.asfortranarray(W, dtype=np.dtype(dtype))
	n, k = Wf.shape
	cols = torch.from_numpy(Wf.T).to("cuda")  # (k, n) C-contiguous == F-ordered (n, k)
	X = torch.empty((n, k), dtype=cols.dtype, device="cuda")
	_ffi.check(_ffi.lib().slq_block_from_columns(cols.data_ptr(), n, k, _code(dtype), X.data_ptr(), _stream_ptr()))
	return X


def block_to_columns(X) -> np.ndarray:
	"""Device probe block (n, k) -> host F-ordered (n, k) array."""
	torch = _torch()
	n, k = int(X.shape[0]), int(X.shape[1])
	cols = torch.empty((k, n), dtype=X.dtype, device=X.device)
	code = _ffi.SLQ_F64 if X.dtype == torch.float64 else _ffi.SLQ_F32
	_ffi.check(_ffi.lib().slq_block_to_columns(X.data_ptr(), n, k, code, cols.data_ptr(), _stream_ptr()))
	return cols.cpu().numpy().T


def stats_update(y, state):
	"""Chan merge of device samples y (m,) into device state (3,) = (count, mean, M2)."""
	_ffi.check(_ffi.lib().slq_stats_update(y.data_ptr(), int(y.numel()), state.data_ptr(), _stream_ptr()))
