"""Small dense helpers (API mirror of src/primate/linalg.py:5-24)."""
import numpy as np


def append_trinv(R_inv: np.ndarray, C: np.ndarray, D_inv: np.ndarray) -> np.ndarray:
	r"""Inverse of the block upper-triangular matrix $\begin{pmatrix} R & C \\ 0 & D \end{pmatrix}$ from $R^{-1}$ (m x m),
	the coupling block $C$ (m x s) and $D^{-1}$ (s x s): $\begin{pmatrix} R^{-1} & -R^{-1} C D^{-1} \\ 0 & D^{-1} \end{pmatrix}$
	-- the batch form of `update_trinv`, used when XTrace appends a block of columns to its QR factorisation."""
	m, s = R_inv.shape[0], D_inv.shape[0]
	out = np.zeros((m + s, m + s))
	out[:m, :m] = R_inv
	out[m:, m:] = D_inv
	if m:
		out[:m, m:] = -(R_inv @ np.reshape(C, (m, s))) @ D_inv
	return out


def update_trinv(B_inv: np.ndarray, b: np.ndarray) -> np.ndarray:
	r"""Inverse of the upper-triangular $[B, b]$ obtained by appending the column `b` (length n + 1) to the n x n
	upper-triangular $B$, given $B^{-1}$ (src/primate/linalg.py:5-24): the one-column case of `append_trinv`."""
	n, m = B_inv.shape
	assert n == m and len(b) == (n + 1), "B must be n x n and `b` must have length `n + 1`"
	b = np.asarray(b).reshape(n + 1)
	assert B_inv.dtype == b.dtype, "dtypes of `B_inv` and `b` did not match."
	return append_trinv(B_inv, b[:n, None], np.array([[1.0 / b[n]]]))
