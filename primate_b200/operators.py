"""Matrix-function operators (API mirror of src/primate/operators.py:15-193) on the device engine.

`MatrixFunction.quad` / `._matvec` keep the reference's host signatures; `quad_block` / `matmat_block` are the
device-resident entries the estimators use (probe blocks in GPU memory in, a few scalars out).  Semantics are the
reference's with a FRESH Lanczos basis per probe (what `primate.lanczos.lanczos` does, lanczos.py:103; the
reference's `MatrixFunction` re-uses a stale `_Q` across probes when orth > 0, which makes its output depend on
probe order -- SURVEY fact 4 -- and is deliberately not reproduced).
"""
from typing import Any, Callable, Optional, Union

import numpy as np
from scipy.sparse.linalg import LinearOperator, aslinearoperator, eigsh
from scipy.sparse.linalg._interface import IdentityOperator

from . import engine, sharding
from .integrate import quadrature  # noqa: F401  (re-exported like the reference module)
from .lanczos import _lanczos, device_operator  # noqa: F401
from .special import param_callable
from .tridiag import eigh_tridiag  # noqa: F401

F64: np.dtype = np.dtype("float64")
BASIS_BUDGET = 40 << 30  # bytes a stored Lanczos basis (deg x n x k) may take before f(A) X goes basis-free
PANEL_BUDGET = 40 << 30  # bytes the five panels of the basis-free engine may take (narrows the block only for n > 3e7)
_FORCE_STORED_BASIS = False  # tests / A-B timing: route every f(A) X through the stored-basis formula
_FORCE_TWOPASS = False  # ... or (orth = 0) through the basis-free two-pass engine


def _op_checks(A: Any) -> list:
	return [hasattr(A, "__matmul__"), hasattr(A, "matmul"), hasattr(A, "dot"), hasattr(A, "matvec")]


def is_valid_operator(A: Union[np.ndarray, LinearOperator]) -> np.dtype:
	"""Asserts `A` is a square float32/float64 operator with a product defined; returns its dtype."""
	assert any(_op_checks(A)), "Invalid operator; must have an overloaded 'matvec' or 'matmul' method"
	assert hasattr(A, "shape") and len(A.shape) >= 2, "Operator must be at least two dimensional."
	assert A.shape[0] == A.shape[1], "This function only works with square, symmetric matrices!"
	f_dtype = A.dtype if hasattr(A, "dtype") else (A @ np.zeros(A.shape[1])).dtype
	assert np.dtype(f_dtype).type in {np.float32, np.float64}, "Only 32- or 64-bit floats are supported."
	return np.dtype(f_dtype)


def is_linear_op(A: Any) -> bool:
	"""Whether `A` offers the interface implicit-matrix algorithms need (a product, a 2-d square shape)."""
	return bool(any(_op_checks(A)) and hasattr(A, "shape") and len(A.shape) >= 2 and A.shape[0] == A.shape[1])


class MatrixFunction(LinearOperator):
	r"""Implicit matrix function $f(A) = U f(\Lambda) U^T$ evaluated by the Lanczos method on the GPU.

	Parameters:
		A: numpy array, sparse matrix, or LinearOperator (uploaded to the device once).
		fun: spectral function: a built-in name ("identity", "abs", "sqrt", "log", "inv", "exp", "smoothstep",
			"numrank"; evaluated on the device) or any vectorised callable (evaluated on the host on the Ritz values).
		deg: degree of the Krylov expansion.
		orth: number of Lanczos vectors to re-orthogonalise against per step.
		dtype: float64 or float32.
		kwargs: parameters of the built-in function (t, a, b, threshold, q).
		basis (keyword, not in the reference): engine behind `M @ X` -- "auto" (default), "stored" (the reference's
			n x deg Lanczos basis), or "twopass" (basis-free: memory independent of deg; orth = 0 only).
		comm (keyword, not in the reference): None (default) keeps block products on this process; "auto" or a
			torch.distributed group splits the columns of every block product / quadratic form over the ranks (the
			operator is replicated per GPU; all ranks must then make the same calls; see `sharding.sharded_columns`).
	"""

	def __init__(self, A: np.ndarray, fun: Optional[Callable] = None, deg: int = 20, orth: int = 3, dtype: np.dtype = F64, **kwargs) -> None:
		assert is_linear_op(A), "Invalid operator `A`; must be dim=2 symmetric operator with defined matvec"
		assert deg >= 2, "Degree must be >= 2"
		comm = kwargs.pop("comm", None)
		self._basis = kwargs.pop("basis", "auto")  # f(A) X engine: "auto" | "stored" | "twopass" (see _use_twopass)
		assert self._basis in ("auto", "stored", "twopass")
		super().__init__(dtype=np.dtype(dtype), shape=tuple(A.shape))  # SciPy >= 1.12 needs this for `M @ v`
		if fun is None or isinstance(fun, str):
			host_fun = param_callable(fun, **dict(kwargs)) if isinstance(fun, str) else (lambda x: x)
			spec = engine.fun_spec(fun, kwargs)  # device-side evaluation of the built-in
		else:
			host_fun, spec = fun, None
		self.fun = host_fun
		self._fun_spec = spec
		self._comm = sharding.resolve_comm(comm)
		self._deg = min(deg, A.shape[0])
		self._rtol = 1e-8
		self._orth = self._deg if orth < 0 or orth > self._deg else orth
		self._A = A
		self._dev = None
		self._nodes = np.zeros(self._deg, dtype=dtype)
		self._weights = np.zeros(self._deg, dtype=dtype)
		self._alpha = np.zeros(self._deg + 1, dtype=dtype)
		self._beta = np.zeros(self._deg + 1, dtype=dtype)

	@property
	def degree(self) -> int:
		return self._deg

	@property
	def fun(self) -> Callable:
		return self._fun

	@fun.setter
	def fun(self, value: Callable) -> None:
		assert callable(value), "Function must be callable."
		out = value(np.ones(self.shape[1]))
		assert isinstance(out, np.ndarray), "Function must return array-like"
		assert out.shape[-1] == self.shape[0], "Last dimension of output must match number of rows."
		self._fun = value
		self._fun_spec = None  # a user callable: evaluated on the host unless __init__ re-instates a built-in id

	def _adjoint(self):
		return self

	@property
	def device_op(self) -> "engine.DeviceOperator":
		if self._dev is None:
			self._dev = device_operator(self._A, self.dtype)
		return self._dev

	# ---- device-resident entries ----------------------------------------------------------------------------
	def quad_block(self, X, keep_tridiag: bool = False):
		"""x_p^T f(A) x_p for every column of the device probe block X (n, k) (clobbered) -> device tensor (k,)."""
		# ncv = max(2, orth) visits exactly the vectors the reference's ncv = deg does (fresh basis), with orth+1 panels
		out = self.device_op.lanczos(X, self._deg, self._rtol, self._orth, max(2, self._orth))
		if self._fun_spec is not None:
			fid, prm = self._fun_spec
			_, _, y, _ = engine.quadrature(out["alpha"], out["beta"], self._deg, fid, prm, scale=out["xnorm2"], want_nodes=False)
		else:
			torch = engine._torch()
			theta, tau, _, _ = engine.quadrature(out["alpha"], out["beta"], self._deg, fun_id=-1)
			th, ta = theta.cpu().numpy(), tau.cpu().numpy()
			vals = np.array([np.sum(self._fun(th[j]) * ta[j], axis=-1) for j in range(th.shape[0])]) * out["xnorm2"].cpu().numpy()
			y = torch.from_numpy(np.asarray(vals, dtype=np.float64)).cuda()
		return (y, out) if keep_tridiag else y

	def quad_range(self, stream, first: int, count: int, kmax: int = 64):
		"""Quadrature samples of probes [first, first+count) of a `ProbeStream` -> device tensor (count,).

		The Lanczos blocks run back to back and write their tridiagonals into one (count, deg+1) array; the Golub-Welsch
		solves of ALL probes then go out as ONE launch (one warp per Jacobi matrix).  Per block of 32 the QL sweep is a
		serial ~1 ms chain on 32 warps of an otherwise idle GPU (6.6% of a C2 step); batched over the range it is noise."""
		if count <= 0 or self._fun_spec is None:
			parts = [self.quad_block(stream.block(k, first + off)) for off, k in _offsets(engine.block_sizes(count, kmax))]
			return engine._torch().cat(parts) if parts else engine._torch().zeros(0, dtype=engine._torch().float64, device="cuda")
		torch = engine._torch()
		tdt = engine._tdtype(self.dtype)
		alpha = torch.empty((count, self._deg + 1), dtype=tdt, device="cuda")
		beta = torch.empty((count, self._deg + 1), dtype=tdt, device="cuda")
		xnorm2 = torch.empty(count, dtype=torch.float64, device="cuda")
		op, ncv = self.device_op, max(2, self._orth)
		for off, k in _offsets(engine.block_sizes(count, kmax)):
			out = {"alpha": alpha[off : off + k], "beta": beta[off : off + k], "xnorm2": xnorm2[off : off + k]}
			op.lanczos(stream.block(k, first + off), self._deg, self._rtol, self._orth, ncv, out=out)
		fid, prm = self._fun_spec
		_, _, y, _ = engine.quadrature(alpha, beta, self._deg, fid, prm, scale=xnorm2, want_nodes=False)
		return y

	def _host_coeffs(self, out):
		"""Krylov coefficients c = Y f(Theta) Y^T e_1 for a Python callable: eigenvectors from the device QL solver, `f`
		evaluated on the host on the (k x deg) Ritz values (operators.py:122-124)."""
		torch = engine._torch()
		_, theta, _, scratch, _ = engine.funm_coeffs(out["alpha"], out["beta"], self._deg, fun_id=-1)
		k, m = theta.shape
		th, sc = theta.cpu().numpy(), scratch.cpu().numpy()
		Zt = sc[: k * m * m].reshape(k, m, m)
		perm = sc[k * m * m :].reshape(k, m).astype(np.int64)
		C = np.empty((k, m))
		for p in range(k):
			Y = Zt[p][perm[p], :].T  # (m, m), columns sorted like theta
			C[p] = Y @ (np.ravel(self._fun(th[p])) * Y[0, :])
		return torch.from_numpy(C).cuda()

	def nodes_range(self, stream, first: int, count: int, kmax: int = 64):
		"""Gauss quadrature rules of probes [first, first + count) of a `ProbeStream`: device tensors theta, tau of shape
		(count, deg) -- Ritz values and weights of every probe, the raw material of spectral densities / moments
		(src/primate/lanczos.py:120-164 and the reference's planned `spectral_density`) -- and ||x||^2 (count,).  The
		Lanczos blocks run back to back; ONE batched Golub-Welsch launch (a warp per Jacobi matrix) follows."""
		torch = engine._torch()
		tdt = engine._tdtype(self.dtype)
		alpha = torch.empty((count, self._deg + 1), dtype=tdt, device="cuda")
		beta = torch.empty((count, self._deg + 1), dtype=tdt, device="cuda")
		xnorm2 = torch.empty(count, dtype=torch.float64, device="cuda")
		op, ncv = self.device_op, max(2, self._orth)
		for off, k in _offsets(engine.block_sizes(count, kmax)):
			out = {"alpha": alpha[off : off + k], "beta": beta[off : off + k], "xnorm2": xnorm2[off : off + k]}
			op.lanczos(stream.block(k, first + off), self._deg, self._rtol, self._orth, ncv, out=out)
		theta, tau, _, info = engine.quadrature(alpha, beta, self._deg, fun_id=-1)
		return theta, tau, xnorm2

	def matmat_block(self, X):
		"""f(A) X for a device probe block X (n, k) -> device block (n, k).

		orth = 0: the basis-free two-pass engine (X is left intact; memory independent of `deg`).  orth > 0: the
		reference's stored-basis formula y = ||x|| Q Y f(Theta) Y^T e_1 (operators.py:102-124; X is clobbered)."""
		torch = engine._torch()
		if self._use_twopass(self.shape[0]):
			if self._fun_spec is not None:
				fid, prm = self._fun_spec
				return self.device_op.funm_matmat(X, self._deg, self._rtol, fid, prm)
			run = self.device_op.lanczos(X.clone(), self._deg, self._rtol, 0, 2)
			return self.device_op.lanczos_replay(X, self._deg, run, self._host_coeffs(run))
		out = self.device_op.lanczos(X, self._deg, self._rtol, self._orth, self._deg, want_basis=True, internal_basis=True)
		if self._fun_spec is not None:
			fid, prm = self._fun_spec
			coeffs, _, _, _, _ = engine.funm_coeffs(out["alpha"], out["beta"], self._deg, fid, prm)
		else:
			coeffs = self._host_coeffs(out)
		Y = torch.empty_like(X)
		from . import _ffi

		_ffi.check(_ffi.lib().slq_basis_combine(out["Q"].data_ptr(), self.shape[0], int(X.shape[1]), self._deg, self.device_op.code,
		                                        coeffs.data_ptr(), out["xnorm2"].data_ptr(), Y.data_ptr(), engine._stream_ptr()))  # fmt: skip
		return self.device_op._from_internal(Y)  # the basis was combined in the operator's row order

	# ---- device-panel API (block drivers that keep their panels on the GPU) -----------------------------------------
	def _use_twopass(self, n: int) -> bool:
		"""f(A) X policy.  Both engines move about the same bytes per column and step (stored basis: CSR/k + 9 panel
		passes; two-pass: 2 CSR/k + 11), so the stored basis -- one Lanczos sweep -- is kept while a basis of at least 4
		columns fits in `BASIS_BUDGET` (measured at config C5, 64 columns: 1.74 s stored at k = 4 vs 1.94 s two-pass at
		k = 32); beyond that, and whenever the caller asks (`basis="twopass"`), the basis-free engine takes over: its
		memory is five panels whatever `deg` is.  orth > 0 needs the stored vectors for the re-projection."""
		if self._orth != 0 or _FORCE_STORED_BASIS or self._basis == "stored":
			return False
		if _FORCE_TWOPASS or self._basis == "twopass":
			return True
		return 4 * self._deg * n * np.dtype(self.dtype).itemsize > BASIS_BUDGET

	def _basis_block_width(self, n: int, budget: Optional[int] = None) -> int:
		"""Widest probe block (power of two <= 32) for f(A) X: the two-pass engine holds five panels whatever `deg` is;
		the stored-basis path is narrowed until its deg x n x k basis stays within the budget."""
		itemsize = np.dtype(self.dtype).itemsize
		if self._use_twopass(n):  # five panels: its own (fixed) allowance, independent of the basis budget that selected it
			budget, per_column = (PANEL_BUDGET if budget is None else budget), 5 * n * itemsize
		else:
			budget, per_column = (BASIS_BUDGET if budget is None else budget), self._deg * n * itemsize
		k = 32
		while k > 1 and k * per_column > budget:
			k >>= 1
		return k

	def _split_columns(self, m: int):
		c = self._comm
		if c.size > 1 and m > 1 and c.backend == "nccl":
			lo, hi = sharding.probe_range(m, c.rank, c.size)
			return c, lo, hi, -(-m // c.size)
		return None, 0, m, m

	def matmat_device(self, Xd):
		"""f(A) X for a device panel Xd (n, m), row-major, left intact -> device panel (n, m).  With `comm` set, every
		rank runs its own column range and the panels are all-gathered over NCCL."""
		torch = engine._torch()
		n, m = int(Xd.shape[0]), int(Xd.shape[1])
		c, lo, hi, per = self._split_columns(m)
		mine = torch.zeros((n, per), dtype=Xd.dtype, device=Xd.device) if c is not None else torch.empty_like(Xd)
		kmax = self._basis_block_width(n)
		done = lo
		for kb in engine.block_sizes(hi - lo, kmax):
			mine[:, done - lo : done - lo + kb] = self.matmat_block(_block_copy(Xd, done, kb))
			done += kb
		if c is None:
			return mine
		allp = c.all_gather_tensor(mine)  # (size, n, per): rank r holds the global columns [r per, r per + per)
		return allp.permute(1, 0, 2).reshape(n, c.size * per)[:, :m].contiguous()

	def quad_device(self, Xd):
		"""x^T f(A) x for every column of the device panel Xd (n, m) (left intact) -> device tensor (m,) float64."""
		torch = engine._torch()
		m = int(Xd.shape[1])
		c, lo, hi, per = self._split_columns(m)
		mine = torch.zeros(per if c is not None else m, dtype=torch.float64, device=Xd.device)
		done = lo
		for kb in engine.block_sizes(hi - lo, 64):
			mine[done - lo : done - lo + kb] = self.quad_block(_block_copy(Xd, done, kb))
			done += kb
		if c is None:
			return mine
		return c.all_gather_tensor(mine).reshape(-1)[:m].contiguous()

	# ---- host-array API of the reference ------------------------------------------------------------------------
	def _matvec(self, x: np.ndarray):
		r"""Estimates $x \mapsto f(A) x$ with a degree-`deg` Krylov expansion (operators.py:102-124)."""
		x = np.asarray(x).astype(self.dtype).reshape(-1, 1)
		Y = self.matmat_block(engine.block_from_columns(x, self.dtype))
		return engine.block_to_columns(Y)

	def _matmat(self, X: np.ndarray):
		X = np.asarray(X).astype(self.dtype)
		if self._comm.size > 1 and X.shape[1] > 1:
			if self._comm.backend == "nccl":
				return self._matmat_sharded(X)
			return sharding.sharded_columns(self._matmat_local, X, self._comm).astype(self.dtype, copy=False)
		return self._matmat_local(X)

	def _matmat_sharded(self, X: np.ndarray):
		"""Columns split over the ranks: each rank uploads and runs its own column range, the result panels
		are all-gathered on the devices (NCCL over NVLink) and downloaded once."""
		import torch

		n, m = X.shape
		c = self._comm
		lo, hi = sharding.probe_range(m, c.rank, c.size)
		per = -(-m // c.size)
		mine = torch.zeros((per, n), dtype=engine._tdtype(self.dtype), device="cuda")
		if hi > lo:
			mine[: hi - lo] = self._matmat_device(X[:, lo:hi])
		allp = c.all_gather_tensor(mine)  # (size, per, n)
		if per * c.size == m:
			return allp.reshape(m, n).cpu().numpy().T
		parts = [allp[r, : b - a] for r in range(c.size) for a, b in [sharding.probe_range(m, r, c.size)] if b > a]
		return torch.cat(parts, 0).cpu().numpy().T

	def _device_blocks(self, X: np.ndarray, kmax: int):
		"""Uploads the host panel X (n, m) ONCE, in the order it has, and yields (first column, k, device probe block)."""
		import torch

		from . import _ffi

		n, m = X.shape
		if not X.flags.c_contiguous and not X.flags.f_contiguous and X.strides[1] == X.itemsize:
			X = np.ascontiguousarray(X)  # a column range of a row-major panel: compact rows, no transpose
		rowmajor = X.flags.c_contiguous
		Xd = torch.from_numpy(X if rowmajor else np.asfortranarray(X).T).to("cuda")  # (n, m) row-major or (m, n) columns
		done = 0
		for kb in engine.block_sizes(m, kmax):
			if rowmajor:
				B = Xd[:, done : done + kb].contiguous()  # already the probe-block layout
			else:
				B = torch.empty((n, kb), dtype=Xd.dtype, device="cuda")
				_ffi.check(_ffi.lib().slq_block_from_columns(Xd[done : done + kb].data_ptr(), n, kb, self.device_op.code, B.data_ptr(),
				                                             engine._stream_ptr()))  # fmt: skip
			yield done, kb, B
			done += kb

	def _matmat_local(self, X: np.ndarray):
		"""One upload of the whole panel, blocked device Lanczos over its columns, one download: no host transposes."""
		return self._matmat_device(X).cpu().numpy().T

	def _matmat_device(self, X: np.ndarray):
		"""f(A) X for a host panel X (n, m) -> device tensor (m, n) holding the result columns."""
		import torch

		from . import _ffi

		n, m = X.shape
		Yt = torch.empty((m, n), dtype=engine._tdtype(self.dtype), device="cuda")  # columns: an F-ordered (n, m) host array
		kmax = self._basis_block_width(n)
		for done, kb, B in self._device_blocks(X, kmax):
			Y = self.matmat_block(B)
			_ffi.check(_ffi.lib().slq_block_to_columns(Y.data_ptr(), n, kb, self.device_op.code, Yt[done : done + kb].data_ptr(),
			                                           engine._stream_ptr()))  # fmt: skip
		return Yt

	def quad(self, x: np.ndarray):
		r"""Estimates the quadratic forms $x \mapsto x^T f(A) x$ of the columns of `x` by Lanczos quadrature
		(operators.py:126-151).  Returns a float64 array with one entry per column."""
		x = np.asarray(x).astype(self.dtype)
		x = np.atleast_2d(x).T if x.ndim == 1 else x
		if self._comm.size > 1 and x.shape[1] > 1:
			return sharding.sharded_columns(lambda b: self._quad_local(b)[np.newaxis, :], x, self._comm).ravel()
		return self._quad_local(x)

	def _quad_local(self, x: np.ndarray):
		m = x.shape[1]
		y = np.zeros(m)
		last = None
		for done, kb, B in self._device_blocks(x, 64):
			yb, last = self.quad_block(B, keep_tridiag=True)
			y[done : done + kb] = yb.cpu().numpy()
		if last is not None:  # like the reference, leave the last probe's tridiagonal entries on the instance
			self._alpha[:] = last["alpha"][-1].cpu().numpy()
			self._beta[:] = last["beta"][-1].cpu().numpy()
		return y


def _block_copy(Xd, first: int, k: int):
	"""Columns [first, first + k) of a row-major device panel as a fresh contiguous probe block (the engine clobbers it)."""
	B = Xd[:, first : first + k]
	return B.clone() if B.is_contiguous() else B.contiguous()


def _offsets(sizes):
	off = 0
	for k in sizes:
		yield off, k
		off += k


def matrix_function(A: LinearOperator, fun: Optional[Callable] = None, v: Optional[np.ndarray] = None, deg: int = 20):
	"""`MatrixFunction(A, fun, deg)`, or its action on `v` when given."""
	M = MatrixFunction(A, fun=fun, deg=deg)
	return M if v is None else M._matvec(v)


class Toeplitz(LinearOperator):
	"""Matrix-free symmetric Toeplitz / circulant operator with an FFT product (host operator: reaches the engine
	through the callback path, like the reference's PyLinearOperator)."""

	def __init__(self, c: np.ndarray, r: Optional[np.ndarray] = None, dtype: np.dtype = F64):
		self.c = np.array(c)
		self.r = np.array(c if r is None else r)
		m = len(self.c)
		super().__init__(dtype=np.dtype(dtype), shape=(m, m))
		self._symbol = np.real(np.fft.fft(np.concatenate((self.c, [0], self.r[:0:-1]))))
		self._pad = np.zeros(2 * m)

	def _matvec(self, x: np.ndarray) -> np.ndarray:
		m = self.shape[0]
		assert len(x) == m, f"Invalid shape of input vector 'x'; must have length {m}"
		self._pad[:m] = np.ravel(x).astype(self.dtype, copy=False)
		y = np.real(np.fft.ifft(self._symbol * np.fft.fft(self._pad)))
		return y[:m].astype(self.dtype)


def normalize_unit(A: LinearOperator, interval: tuple = (-1, 1)) -> LinearOperator:
	"""Shifts and scales `A` so that its spectrum lies in [-1, 1] (uses the largest-magnitude eigenvalue)."""
	A = aslinearoperator(A) if not isinstance(A, LinearOperator) else A
	assert isinstance(A, LinearOperator), "A must be a linear operator"
	lam = eigsh(A, k=1, which="LM", return_eigenvectors=False).item()
	return (A + lam * IdentityOperator(A.shape)) / (2 * lam)
