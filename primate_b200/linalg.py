"""Small dense helpers (API mirror of src/primate/linalg.py:5-24)."""
import numpy as np


def update_trinv(B_inv: np.ndarray, b: np.ndarray) -> np.ndarray:
	r"""Inverse of the upper-triangular matrix $[B, b]$ obtained by appending the column `b` (length n+1) to $B$,
	given $B^{-1}$ (n x n): the new last column is $-B^{-1} b_{:n} / b_n$ with $1 / b_n$ on the diagonal."""
	n, m = B_inv.shape
	assert n == m and len(b) == (n + 1), "B must be n x n and `b` must have length `n + 1`"
	b = np.atleast_2d(b).reshape((n + 1, 1))
	assert B_inv.dtype == b.dtype, "dtypes of `B_inv` and `b` did not match."
	out = np.zeros(shape=(n + 1, n + 1))
	out[:n, :n] = B_inv
	pivot = np.reciprocal(b[-1]).item()
	out[n, n] = pivot
	out[:n, [-1]] = pivot * ((-B_inv) @ b[:-1])
	return out
