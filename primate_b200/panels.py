"""Tall-skinny panel algebra behind the block drivers (hutch++, XDiag, XTrace).

The drivers in `trace.py` / `diagonal.py` are written ONCE against the small interface below and run on either backend:

* `DevicePanels` -- n x m panels are row-major torch tensors that never leave the GPU; every pass over them is a libslq
  kernel (csrc/slq_panels.cu: slq_panel_gram / _matmul / _rowdot / _coldot), products with the operator are blocked
  device Lanczos (`MatrixFunction.matmat_device`), probes come from the device generators.  Thin QR is CholeskyQR2; only
  m x m matrices (Gram matrices, Cholesky factors, triangular inverses) visit the host.  With a `comm` (one process
  per GPU) the Gram-type reductions run over each rank's slice of the rows and are summed by ONE all-reduce of m^2
  doubles; the operator products are split by columns (`MatrixFunction.matmat_device`).
* `HostPanels` -- NumPy arrays, for operators that are not a `MatrixFunction` (plain matrices, user LinearOperators): the
  reference's setting (np.linalg.qr, `A @ block`).

Interface (X, Y panels with m columns; S an m x m' host ndarray):
  probes(first, count) -> panel        product(X) -> A X          quad(X) -> host (m,) of x^T A x
  gram(X, Y) -> host X^T Y             matmul(X, S) -> X S         axpy_matmul(Y, X, S): Y -= X S (in place)
  rowdot(X, Y) -> n-vector             coldot(X, Y) -> host (m,)   qr(Y) -> (Q panel, R host), R upper triangular
  cat(panels) / cols(X, a, b) / to_host(vector)
"""
from __future__ import annotations

from typing import Optional

import numpy as np


def chol_upper(G: np.ndarray) -> np.ndarray:
	"""Upper-triangular R with R^T R = G for a Gram matrix; a failed factorisation (numerically rank-deficient panel) is
	retried with the shift of shifted CholeskyQR (Fukaya et al. 2020): s = 11 (m + 1) m u ||G||_2."""
	G = (G + G.T) / 2
	try:
		return np.linalg.cholesky(G).T
	except np.linalg.LinAlgError:
		m = G.shape[0]
		shift = 11.0 * (m + 1) * m * np.finfo(np.float64).eps * np.linalg.norm(G, 2)
		return np.linalg.cholesky(G + shift * np.eye(m)).T


def upper_inverse(R: np.ndarray) -> np.ndarray:
	from scipy.linalg import solve_triangular

	return solve_triangular(R, np.eye(R.shape[0]), lower=False)


class HostPanels:
	"""NumPy backend: `A` is anything with `@`; probes from a host sampler `draw(size=(n, count))`."""

	on_device = False

	def __init__(self, A, draw, dtype):
		self.A, self.draw, self.dtype = A, draw, np.dtype(dtype)
		self.n = int(A.shape[0])

	def probes(self, first: int, count: int):
		return self.draw(size=(self.n, count)).astype(self.dtype)

	def product(self, X):
		return self.A @ X

	def quad(self, X):
		if hasattr(self.A, "quad"):
			return np.ravel(self.A.quad(X))
		return np.einsum("ij,ij->j", X, self.A @ X)

	def gram(self, X, Y):
		return X.T @ Y

	def matmul(self, X, S):
		return X @ S

	def axpy_matmul(self, Y, X, S):
		Y -= X @ S
		return Y

	def rowdot(self, X, Y):
		return np.einsum("ij,ij->i", X, Y)

	def coldot(self, X, Y):
		return np.einsum("ij,ij->j", X, Y)

	def qr(self, Y):
		return np.linalg.qr(Y, mode="reduced")

	def cat(self, parts):
		return np.concatenate(parts, axis=1) if len(parts) > 1 else parts[0]

	def empty(self, m: int):
		return np.zeros((self.n, m), dtype=self.dtype)

	def assign(self, X, a: int, block):
		X[:, a : a + block.shape[1]] = block

	def cols(self, X, a: int, b: int):
		return X[:, a:b]

	def to_host(self, v):
		return np.asarray(v, dtype=np.float64)

	def row_slice(self):
		return 0, self.n

	def rows(self, X, lo: int, hi: int):
		return X[lo:hi]

	def gather_rows(self, v_local):
		return v_local


class DevicePanels:
	"""libslq backend for a `MatrixFunction` M and a device `ProbeStream`."""

	on_device = True

	def __init__(self, M, stream, comm=None):
		from . import _ffi, engine

		self.M, self.stream = M, stream
		self.n = int(M.shape[0])
		self.dtype = np.dtype(M.dtype)
		self.code = engine._code(self.dtype)
		self.torch = engine._torch()
		self._ffi, self._engine = _ffi, engine
		c = comm if comm is not None else getattr(M, "_comm", None)
		self.comm = c if (c is not None and c.size > 1 and c.backend == "nccl") else None
		self._ws = None

	# ---- plumbing -----------------------------------------------------------------------------------------------------
	def _scratch(self, nbytes: int):
		if self._ws is None or self._ws.numel() < nbytes:
			self._ws = self.torch.empty(max(int(nbytes), 1), dtype=self.torch.uint8, device="cuda")
		return self._ws

	@staticmethod
	def _ld(X) -> int:
		assert X.dim() == 2 and X.stride(1) == 1, "panels are row-major (unit column stride)"
		return int(X.stride(0))

	def _rows(self):
		"""This rank's slice of the panel rows for the reductions (everything without a comm)."""
		if self.comm is None:
			return 0, self.n
		per = -(-self.n // self.comm.size)
		lo = min(self.comm.rank * per, self.n)
		return lo, min(lo + per, self.n)

	def _allreduce(self, t):
		if self.comm is not None:
			self.comm.all_reduce_sum(t)
		return t

	# ---- interface ----------------------------------------------------------------------------------------------------
	def probes(self, first: int, count: int):
		parts, off = [], 0
		for k in self._engine.block_sizes(count, 64):
			parts.append(self.stream.block(k, first + off))
			off += k
		return self.cat(parts)

	def product(self, X):
		return self.M.matmat_device(X if X.is_contiguous() else X.contiguous())

	def quad(self, X):
		return self.M.quad_device(X if X.is_contiguous() else X.contiguous()).double().cpu().numpy()

	def gram(self, X, Y) -> np.ndarray:
		L = self._ffi.lib()
		a, b = int(X.shape[1]), int(Y.shape[1])
		G = self.torch.empty((a, b), dtype=self.torch.float64, device="cuda")
		need = L.slq_panel_gram_workspace_bytes(a, b)
		ws = self._scratch(need)
		lo, hi = self._rows()
		self._ffi.check(L.slq_panel_gram(X.data_ptr(), a, self._ld(X), Y.data_ptr(), b, self._ld(Y), lo, hi, self.code, G.data_ptr(),
		                                 ws.data_ptr(), ws.numel(), self._engine._stream_ptr()))  # fmt: skip
		return self._allreduce(G).cpu().numpy()

	def _matmul(self, X, S: np.ndarray, alpha: float, beta: float, out):
		L = self._ffi.lib()
		S = np.ascontiguousarray(S, dtype=np.float64)
		a, b = S.shape
		assert int(X.shape[1]) == a
		Sd = self.torch.from_numpy(S).cuda()
		rows = int(X.shape[0])  # the full panel, or a rank's row slice of it
		if out is None:
			out = self.torch.empty((rows, b), dtype=X.dtype, device="cuda")
		self._ffi.check(L.slq_panel_matmul(X.data_ptr(), a, self._ld(X), Sd.data_ptr(), b, float(alpha), float(beta), out.data_ptr(),
		                                   self._ld(out), rows, self.code, self._engine._stream_ptr()))  # fmt: skip
		return out

	def matmul(self, X, S):
		return self._matmul(X, S, 1.0, 0.0, None)

	def axpy_matmul(self, Y, X, S):
		return self._matmul(X, S, -1.0, 1.0, Y)

	def rowdot(self, X, Y):
		L = self._ffi.lib()
		rows = int(X.shape[0])
		out = self.torch.empty(rows, dtype=self.torch.float64, device="cuda")
		self._ffi.check(L.slq_panel_rowdot(X.data_ptr(), self._ld(X), Y.data_ptr(), self._ld(Y), int(X.shape[1]), rows, self.code,
		                                   out.data_ptr(), self._engine._stream_ptr()))  # fmt: skip
		return out

	def coldot(self, X, Y) -> np.ndarray:
		L = self._ffi.lib()
		m = int(X.shape[1])
		out = self.torch.empty(m, dtype=self.torch.float64, device="cuda")
		ws = self._scratch(L.slq_panel_coldot_workspace_bytes(m))
		lo, hi = self._rows()
		self._ffi.check(L.slq_panel_coldot(X.data_ptr(), self._ld(X), Y.data_ptr(), self._ld(Y), m, lo, hi, self.code, out.data_ptr(),
		                                   ws.data_ptr(), ws.numel(), self._engine._stream_ptr()))  # fmt: skip
		return self._allreduce(out).cpu().numpy()

	def qr(self, Y):
		"""CholeskyQR2: two rounds of (Gram matrix -> Cholesky on the host -> Y R^-1 as a streaming kernel); the second
		round restores orthogonality to machine precision for cond(Y) < ~1e7 (Yamamoto et al. 2015)."""
		R1 = chol_upper(self.gram(Y, Y))
		Q1 = self.matmul(Y, upper_inverse(R1))
		R2 = chol_upper(self.gram(Q1, Q1))
		Q = self.matmul(Q1, upper_inverse(R2))
		return Q, R2 @ R1

	def cat(self, parts):
		return self.torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]

	def empty(self, m: int):
		return self.torch.zeros((self.n, m), dtype=self._engine._tdtype(self.dtype), device="cuda")

	def assign(self, X, a: int, block):
		X[:, a : a + int(block.shape[1])] = block

	def cols(self, X, a: int, b: int):
		return X[:, a:b]

	def to_host(self, v):
		return v.double().cpu().numpy() if hasattr(v, "cpu") else np.asarray(v, dtype=np.float64)

	def row_slice(self):
		"""Rows whose entries of an n-vector result this rank computes (all of them without a comm): everything after the
		thin QR in XDiag is row-local, so the replicated panels are only ever multiplied on a 1/ranks slice of their rows."""
		return self._rows()

	def rows(self, X, lo: int, hi: int):
		return X[lo:hi]

	def gather_rows(self, v_local):
		"""Concatenates the ranks' row slices of an n-vector (one all-gather of n doubles)."""
		if self.comm is None:
			return v_local
		per = -(-self.n // self.comm.size)
		pad = self.torch.zeros(per, dtype=self.torch.float64, device="cuda")
		pad[: v_local.numel()] = v_local
		return self.comm.all_gather_tensor(pad).reshape(-1)[: self.n]


def block_orthogonalize(P, Q: Optional[object], Ynew):
	"""Appends the columns of Ynew to an orthonormal panel Q: two classical Gram-Schmidt sweeps against Q, then a thin QR
	of the remainder.  Returns (Q_new block, C = Q^T Ynew (m x ns) or None, R_new (ns x ns)): the QR factor of
	[Y_old, Ynew] is [[R, C], [0, R_new]] -- the batch form of the reference's column-by-column `qr_insert`
	(src/primate/trace.py:298-300)."""
	C = None
	if Q is not None and int(Q.shape[1]) > 0:
		C1 = P.gram(Q, Ynew)
		P.axpy_matmul(Ynew, Q, C1)
		C2 = P.gram(Q, Ynew)
		P.axpy_matmul(Ynew, Q, C2)
		C = C1 + C2
	Qn, Rn = P.qr(Ynew)
	return Qn, C, Rn
