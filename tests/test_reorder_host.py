"""Host-side operator ingestion entry points of libslq (no GPU needed): locality ordering and CSR permutation."""
import numpy as np
import scipy.sparse as sp

from .conftest import laplacian_2d


def test_cluster_order_is_a_locality_permutation():
	"""slq_csr_cluster_order / slq_csr_permute (host entry points): a permutation; P A P^T exact; more tile-local refs."""
	from primate_b200 import _ffi

	L = _ffi.lib()
	A = laplacian_2d(90)
	n = A.shape[0]
	ip, ix, va = A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data.astype(np.float64)
	perm = np.empty(n, dtype=np.int32)
	_ffi.check(L.slq_csr_cluster_order(n, ip.ctypes.data, ix.ctypes.data, 120, perm.ctypes.data))
	assert np.array_equal(np.sort(perm), np.arange(n))
	ip2, ix2, va2 = np.empty_like(ip), np.empty_like(ix), np.empty_like(va)
	_ffi.check(L.slq_csr_permute(n, ip.ctypes.data, ix.ctypes.data, va.ctypes.data, 1, perm.ctypes.data, ip2.ctypes.data, ix2.ctypes.data, va2.ctypes.data))
	B = sp.csr_matrix((va2, ix2, ip2), shape=A.shape)
	ref = A[perm][:, perm].tocsr()
	ref.sort_indices()
	assert np.array_equal(B.indptr, ref.indptr) and np.array_equal(B.indices, ref.indices) and np.array_equal(B.data, ref.data)
	local = lambda M: np.mean(M.tocoo().row // 120 == M.tocoo().col // 120)
	assert local(B) > local(A) + 0.15


def test_cluster_order_handles_empty_rows_and_disconnected_parts():
	from primate_b200 import _ffi

	L = _ffi.lib()
	A = sp.block_diag([laplacian_2d(7), sp.csr_matrix((5, 5)), laplacian_2d(4)]).tocsr()
	n = A.shape[0]
	ip, ix = A.indptr.astype(np.int32), A.indices.astype(np.int32)
	perm = np.empty(n, dtype=np.int32)
	_ffi.check(L.slq_csr_cluster_order(n, ip.ctypes.data, ix.ctypes.data, 8, perm.ctypes.data))
	assert np.array_equal(np.sort(perm), np.arange(n))
