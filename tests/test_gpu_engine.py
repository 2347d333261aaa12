"""GPU parity of the C-ABI engine (libslq.so) against the golden vectors from the unmodified reference and
against the CPU oracle on seeded inputs.  Bit-exact for probe signs / raw PCG64; 1e-10 rel for fp64 alpha/beta
(north_star), 1e-4 for fp32; 1e-8 rel for quadrature values."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import laplacian_2d, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
	import torch

	assert torch.cuda.is_available(), "GPU tests need a CUDA device"
	from primate_b200 import engine

	return engine


def _L30(golden):
	return sp.csr_matrix((golden["L30_data"], golden["L30_indices"], golden["L30_indptr"]), shape=(900, 900))


def _run_lanczos(eng, A, V, deg, orth=0, ncv=2, rtol=1e-8, dtype=np.float64, basis=False):
	op = eng.DeviceOperator(A, dtype=dtype)
	V = np.asarray(V, dtype=dtype)
	V = V[:, None] if V.ndim == 1 else V
	X = eng.block_from_columns(V, dtype)
	out = op.lanczos(X, deg, rtol, orth, ncv, want_basis=basis)
	res = {k: (v.cpu().numpy() if v is not None else None) for k, v in out.items()}
	return res


def test_device_is_blackwell(eng):
	import ctypes

	from primate_b200 import _ffi

	sm, ma, mi = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
	_ffi.check(_ffi.lib().slq_device_info(ctypes.byref(sm), ctypes.byref(ma), ctypes.byref(mi)))
	assert ma.value == 10 and sm.value > 0


def test_pcg64_raw_bit_exact(eng, oracle, golden):
	st, inc = oracle.pcg64_state(0)
	assert np.array_equal(eng.pcg64_raw(st, inc, 0, 64), golden["raw_seed0_first64"])
	bg = np.random.PCG64(0)
	bg.advance(10**12 + 7)
	assert np.array_equal(eng.pcg64_raw(st, inc, 10**12 + 7, 1000), bg.random_raw(1000))


@pytest.mark.parametrize("seed,shape,key", [(123, (1000, 5), "signs_seed123_1000x5"), (99, (257, 33), "signs_seed99_257x33")])
def test_rademacher_bit_exact_golden(eng, oracle, golden, seed, shape, key):
	N, k = shape
	st, inc = oracle.pcg64_state(seed)
	G = golden[key].astype(np.float64) * 2 - 1
	done = 0
	for kb in eng.block_sizes(k, 32):
		X = eng.rademacher_block(st, inc, done, kb, N, np.float64).cpu().numpy()
		assert np.array_equal(X, G[:, done : done + kb])
		done += kb
	X32 = eng.rademacher_block(st, inc, 1, 4, N, np.float32).cpu().numpy()
	assert X32.dtype == np.float32 and np.array_equal(X32, G[:, 1:5])


def test_rademacher_large_offsets(eng, oracle):
	"""probe index far into the stream (rank sharding): first_probe * n beyond 2^40."""
	n, first = 100_003, 50_000_017
	st, inc = oracle.pcg64_state(7)
	X = eng.rademacher_block(st, inc, first, 2, n, np.float64).cpu().numpy()
	for p in range(2):
		ref = oracle.c_rademacher(st, inc, (first + p) * n, n)
		assert np.array_equal(X[:, p], ref)


@pytest.mark.parametrize("k", [1, 2, 4, 8, 16, 32, 64])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_spmm_vs_scipy(eng, k, dtype):
	rng = np.random.default_rng(k)
	A = laplacian_2d(37, dtype)
	X = rng.standard_normal((A.shape[0], k)).astype(dtype)
	op = eng.DeviceOperator(A, dtype=dtype)
	Y = op.matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = A @ X
	assert np.allclose(Y, ref, rtol=1e-13 if dtype == np.float64 else 1e-5, atol=1e-13 if dtype == np.float64 else 1e-5)


def test_spmm_random_sparse_and_dense(eng):
	rng = np.random.default_rng(3)
	S = sp.random(501, 501, density=0.02, random_state=5, format="csr")
	S = (S + S.T).tocsr()
	X = rng.standard_normal((501, 8))
	Y = eng.DeviceOperator(S).matmat(eng.block_from_columns(X, np.float64)).cpu().numpy()
	assert np.allclose(Y, S @ X, rtol=1e-12, atol=1e-12)
	D = rng.standard_normal((333, 333))
	D = D + D.T
	for k in (1, 4, 32):
		Xk = rng.standard_normal((333, k))
		Y = eng.DeviceOperator(D).matmat(eng.block_from_columns(Xk, np.float64)).cpu().numpy()
		assert np.allclose(Y, D @ Xk, rtol=1e-11, atol=1e-11)


@pytest.mark.parametrize("k", [4, 8, 16, 32, 64])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_dense_matmat_tiled_path(eng, k, dtype):
	"""Dense symmetric operator with n a multiple of 128 (bulk-copy staged kernel for k >= 8) against NumPy."""
	rng = np.random.default_rng(100 + k)
	n = 384
	D = rng.standard_normal((n, n))
	D = ((D + D.T) / 2).astype(dtype)
	X = rng.standard_normal((n, k)).astype(dtype)
	Y = eng.DeviceOperator(D, dtype=dtype).matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = D.astype(np.float64) @ X.astype(np.float64)
	assert np.allclose(Y, ref, rtol=1e-12 if dtype == np.float64 else 2e-4, atol=1e-11 if dtype == np.float64 else 2e-3)


@pytest.mark.parametrize("rpt", [8, 7, 6])
@pytest.mark.parametrize("n,k,dtype", [(1232, 32, np.float64), (400, 8, np.float64), (2000, 64, np.float32), (16, 16, np.float64), (4096, 32, np.float64)])
def test_dense_matmat_tile_heights(eng, monkeypatch, rpt, n, k, dtype):
	"""Every compiled CTA height (128, 112, 96 rows) of the warp-specialised dense kernel, with a short last row tile and
	more chunks than pipeline stages, against NumPy (SLQ_DENSE_RPT is read at every launch)."""
	monkeypatch.setenv("SLQ_DENSE_RPT", str(rpt))
	rng = np.random.default_rng(n + k + rpt)
	D = rng.standard_normal((n, n))
	D = ((D + D.T) / 2).astype(dtype)
	X = rng.standard_normal((n, k)).astype(dtype)
	Y = eng.DeviceOperator(D, dtype=dtype).matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = D.astype(np.float64) @ X.astype(np.float64)
	assert np.allclose(Y, ref, rtol=1e-12 if dtype == np.float64 else 2e-4, atol=1e-10 if dtype == np.float64 else 5e-3)


@pytest.mark.parametrize("n", [17, 100, 1000, 1231])
@pytest.mark.parametrize("k,dtype", [(64, np.float64), (8, np.float64), (32, np.float32), (1, np.float64)])
def test_dense_matmat_any_order_through_padded_tiles(eng, n, k, dtype):
	"""Orders that are not a multiple of 16 are zero-padded on the device at ingestion and take the bulk-copy kernels
	(DMMA tiles in fp64): short last X slab, stores stop at n, panels keep n rows."""
	rng = np.random.default_rng(n + k)
	D = rng.standard_normal((n, n))
	D = ((D + D.T) / 2).astype(dtype)
	X = rng.standard_normal((n, k)).astype(dtype)
	op = eng.DeviceOperator(D, dtype=dtype)
	Y = op.matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = D.astype(np.float64) @ X.astype(np.float64)
	assert Y.shape == (n, k) and np.all(np.isfinite(Y))
	assert np.allclose(Y, ref, rtol=1e-12 if dtype == np.float64 else 2e-4, atol=1e-10 if dtype == np.float64 else 5e-3)


def test_lanczos_dense_tiled_vs_oracle(eng, oracle):
	"""alpha/beta on a dense SPD matrix of order 256 (tiled dense kernel, k = 8) against the compiled reference."""
	rng = np.random.default_rng(8)
	B = rng.standard_normal((256, 256))
	A = B @ B.T / 256 + 0.5 * np.eye(256)
	st, inc = oracle.pcg64_state(21)
	X = eng.rademacher_block(st, inc, 0, 8, 256, np.float64)
	W = X.cpu().numpy().copy()
	out = eng.DeviceOperator(A).lanczos(X, 15, 1e-8, 0, 2)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	for p in (0, 3, 7):
		ra, rb, _ = oracle.ref_lanczos(A, W[:, p], 15, 1e-8, 0)
		assert rel_err(al[p, :15], ra[:15]) < 1e-10 and rel_err(be[p, 1:15], rb[1:15]) < 1e-10


def _ragged_matrix(n, seed, dtype=np.float64, heavy_rows=(0, 7, 1500), heavy_nnz=6000):
	"""Symmetric CSR with empty rows, a few very long rows (several stage-capacity chunks in the TMA step-A kernel),
	and n / nnz that are not multiples of any tile size."""
	rng = np.random.default_rng(seed)
	S = sp.random(n, n, density=3.0 / n, random_state=seed, format="lil", dtype=np.float64)
	for r in heavy_rows:
		if r < n:
			cols = rng.choice(n, size=min(heavy_nnz, n), replace=False)
			S[r, cols] = rng.standard_normal(cols.size)
	S = sp.csr_matrix(S)
	S = (S + S.T).tolil()
	for r in (3, 4, 5, n - 1):  # empty rows / columns (also the last row)
		S[r, :] = 0
		S[:, r] = 0
	S = sp.csr_matrix(S).astype(dtype)
	S.eliminate_zeros()
	S.sort_indices()
	return S


@pytest.mark.parametrize("k", [8, 16, 32, 64])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_spmm_ragged_long_rows(eng, k, dtype):
	"""Empty rows, rows of ~6000 non-zeros (multi-chunk tiles), ragged n: exercised through slq_op_matmat."""
	A = _ragged_matrix(7013, 17, dtype)
	assert np.diff(A.indptr).max() > 5000 and np.diff(A.indptr).min() == 0
	rng = np.random.default_rng(k)
	X = rng.standard_normal((A.shape[0], k)).astype(dtype)
	Y = eng.DeviceOperator(A, dtype=dtype).matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = (A.astype(np.float64) @ X.astype(np.float64))
	scale = np.abs(A.astype(np.float64)) @ np.abs(X.astype(np.float64)) + 1.0
	assert np.max(np.abs(Y - ref) / scale) < (1e-13 if dtype == np.float64 else 2e-5)


@pytest.mark.parametrize("dtype,k", [(np.float64, 32), (np.float32, 32), (np.float64, 8), (np.float32, 64)])
def test_spmm_power_law_graph_laplacian(eng, dtype, k):
	"""Graph Laplacian with a power-law degree sequence (hubs of ~3000, many rows of 1-100 non-zeros, isolated vertices):
	the non-zero-balanced step-A variant, rows split over slices and chunks in every possible way."""
	rng = np.random.default_rng(12)
	n = 20011
	w = 1.0 / np.arange(1, n + 1) ** 0.9
	src = rng.choice(n, size=300000, p=w / w.sum())
	dst = rng.integers(0, n, size=300000)
	keep = src != dst
	W = sp.coo_matrix((np.ones(keep.sum()), (src[keep], dst[keep])), shape=(n, n)).tocsr()
	W = ((W + W.T) > 0).astype(np.float64)
	L = (sp.diags(np.asarray(W.sum(axis=1)).ravel()) - W).tocsr().astype(dtype)
	L.sort_indices()
	assert np.diff(L.indptr).max() > 2000
	X = rng.standard_normal((n, k)).astype(dtype)
	Y = eng.DeviceOperator(L, dtype=dtype, reorder=False).matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
	ref = L.astype(np.float64) @ X.astype(np.float64)
	scale = np.abs(L.astype(np.float64)) @ np.abs(X.astype(np.float64)) + 1.0
	assert np.max(np.abs(Y - ref) / scale) < (1e-13 if dtype == np.float64 else 2e-5)
	for order in (True, "balance"):  # locality clusters; rows dealt to the tiles by length (the order for such graphs)
		op = eng.DeviceOperator(L, dtype=dtype, reorder=order)
		assert op.perm is not None and np.array_equal(np.sort(op.perm.cpu().numpy()), np.arange(n))
		Y2 = op.matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
		assert np.max(np.abs(Y2 - ref) / scale) < (1e-13 if dtype == np.float64 else 2e-5)


@pytest.mark.parametrize("seed", range(8))
def test_spmm_irregular_random_shapes(eng, seed):
	"""Randomised shapes for the non-zero-balanced step-A variant: sizes around the 120-row tile and the chunk capacity,
	hubs in the first, a middle and the last (partial) tile, empty rows, every block width and both precisions; SpMM
	and three Lanczos steps (alpha through the fused epilogue) against SciPy / the natural-order operator."""
	rng = np.random.default_rng(100 + seed)
	n = int(rng.choice([97, 121, 240, 1000, 2399, 4097, 7919]))
	k = int(rng.choice([8, 16, 32, 64]))
	dtype = np.float64 if seed % 2 == 0 else np.float32
	base = sp.random(n, n, density=min(0.5, 4.0 / n), random_state=int(rng.integers(1 << 30)), format="lil")
	for hub in {0, n // 2, n - 1}:  # rows far longer than 64 non-zeros, up to completely full
		cols = rng.choice(n, size=int(rng.integers(min(n, 70), n + 1)), replace=False)
		base[hub, cols] = rng.standard_normal(len(cols))
	A = base.tocsr()
	A = (A + A.T).tocsr().astype(dtype)
	A.sort_indices()
	assert np.diff(A.indptr).max() > 64
	X = rng.standard_normal((n, k)).astype(dtype)
	ref = A.astype(np.float64) @ X.astype(np.float64)
	scale = np.abs(A.astype(np.float64)) @ np.abs(X.astype(np.float64)) + 1.0
	tol = 1e-13 if dtype == np.float64 else 2e-5
	ops = [eng.DeviceOperator(A, dtype=dtype, reorder=r) for r in (False, "balance")]
	for op in ops:
		Y = op.matmat(eng.block_from_columns(X, dtype)).cpu().numpy()
		assert np.max(np.abs(Y - ref) / scale) < tol
	outs = [op.lanczos(eng.block_from_columns(X, dtype), 3, 1e-8, 0, 2) for op in ops]
	a0, a1 = (o["alpha"].cpu().numpy()[:, :3] for o in outs)
	assert rel_err(a1, a0) < (1e-10 if dtype == np.float64 else 1e-3)
	# alpha_1 = x^T A x / x^T x per probe
	want = np.einsum("ij,ij->j", X.astype(np.float64), ref) / np.einsum("ij,ij->j", X.astype(np.float64), X.astype(np.float64))
	assert np.max(np.abs(a0[:, 0] - want)) < (1e-10 if dtype == np.float64 else 1e-3) * max(1.0, np.max(np.abs(want)))


def test_auto_order_balances_power_law_rows(eng):
	"""`reorder="auto"` on a large operator with hub rows picks the row-balancing order (every 120-row tile then holds
	about the same number of non-zeros); SpMM and Lanczos coefficients are those of the caller's order."""
	rng = np.random.default_rng(3)
	n = 70001
	w = 1.0 / np.arange(1, n + 1) ** 0.95
	src = rng.choice(n, size=700000, p=w / w.sum())
	dst = rng.integers(0, n, size=700000)
	keep = src != dst
	W = sp.coo_matrix((np.ones(keep.sum()), (src[keep], dst[keep])), shape=(n, n)).tocsr()
	W = ((W + W.T) > 0).astype(np.float64)
	L = (sp.diags(np.asarray(W.sum(axis=1)).ravel()) - W).tocsr()
	L.sort_indices()
	deg = np.diff(L.indptr)
	assert deg.max() > max(1024, 64 * L.nnz // n)
	op = eng.DeviceOperator(L)  # reorder="auto"
	perm = op.perm.cpu().numpy()
	assert np.array_equal(np.sort(perm), np.arange(n))
	per_tile = np.add.reduceat(deg[perm], np.arange(0, n, 120))[:-1]
	nat_tile = np.add.reduceat(deg, np.arange(0, n, 120))[:-1]
	assert per_tile.max() - np.median(per_tile) <= deg.max() and nat_tile.max() > 3 * per_tile.max() - 2 * deg.max()
	X = rng.standard_normal((n, 32))
	Y = op.matmat(eng.block_from_columns(X, np.float64)).cpu().numpy()
	scale = np.abs(L) @ np.abs(X) + 1.0
	assert np.max(np.abs(Y - L @ X) / scale) < 1e-13
	nat = eng.DeviceOperator(L, reorder=False)
	assert nat.perm is None
	o1 = op.lanczos(eng.block_from_columns(X, np.float64), 10, 1e-8, 0, 2)
	o2 = nat.lanczos(eng.block_from_columns(X, np.float64), 10, 1e-8, 0, 2)
	assert rel_err(o1["alpha"].cpu().numpy()[:, :10], o2["alpha"].cpu().numpy()[:, :10]) < 1e-10
	assert rel_err(o1["beta"].cpu().numpy()[:, 1:10], o2["beta"].cpu().numpy()[:, 1:10]) < 1e-10


@pytest.mark.parametrize("k", [16, 32, 64])
def test_lanczos_ragged_vs_oracle(eng, oracle, k):
	"""alpha/beta of the blocked device Lanczos on the ragged matrix against the compiled reference, per probe."""
	A = _ragged_matrix(3001, 5, heavy_rows=(0, 11), heavy_nnz=2900)
	n = A.shape[0]
	st, inc = oracle.pcg64_state(3)
	X = eng.rademacher_block(st, inc, 0, k, n, np.float64)
	W = X.cpu().numpy().copy()
	out = eng.DeviceOperator(A).lanczos(X, 12, 1e-8, 0, 2)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	for p in (0, k // 2, k - 1):
		ra, rb, _ = oracle.ref_lanczos(A, W[:, p], 12, 1e-8, 0)
		assert rel_err(al[p, :12], ra[:12]) < 1e-10 and rel_err(be[p, 1:12], rb[1:12]) < 1e-10


@pytest.mark.parametrize("k", [1, 8, 32])
def test_reordered_operator_matches_natural_order(eng, golden, k):
	"""A DeviceOperator with a locality permutation gives the same SpMM, alpha/beta and basis in the caller's row order."""
	A = _L30(golden)
	n = A.shape[0]
	rng = np.random.default_rng(k)
	V = rng.standard_normal((n, k))
	op_r = eng.DeviceOperator(A, reorder=64)
	assert op_r.perm is not None and not np.array_equal(op_r.perm.cpu().numpy(), np.arange(n))
	Y = op_r.matmat(eng.block_from_columns(V, np.float64)).cpu().numpy()
	assert np.allclose(Y, A @ V, rtol=1e-13, atol=1e-13)
	op_n = eng.DeviceOperator(A, reorder=False)
	o_r = op_r.lanczos(eng.block_from_columns(V, np.float64), 20, 1e-8, 3, 20, want_basis=True)
	o_n = op_n.lanczos(eng.block_from_columns(V, np.float64), 20, 1e-8, 3, 20, want_basis=True)
	assert rel_err(o_r["alpha"].cpu().numpy(), o_n["alpha"].cpu().numpy()) < 1e-11
	assert rel_err(o_r["beta"].cpu().numpy(), o_n["beta"].cpu().numpy()) < 1e-11
	assert np.allclose(o_r["Q"].cpu().numpy(), o_n["Q"].cpu().numpy(), atol=1e-9)


def test_reordered_quad_and_matvec_golden(eng, golden):
	"""MatrixFunction.quad / _matvec through a reordered operator against the reference's values on the same inputs."""
	from primate_b200.operators import MatrixFunction

	A = _L30(golden)
	M = MatrixFunction(A, "exp", deg=20, orth=0)
	M._dev = eng.DeviceOperator(A, reorder=64)
	Mn = MatrixFunction(A, "exp", deg=20, orth=0)
	Mn._dev = eng.DeviceOperator(A, reorder=False)
	rng = np.random.default_rng(0)
	V = rng.choice([-1.0, 1.0], size=(A.shape[0], 5))
	assert rel_err(M.quad(V), Mn.quad(V)) < 1e-11
	y, yn = M._matvec(V[:, 0]), Mn._matvec(V[:, 0])
	assert np.allclose(y, yn, rtol=1e-9, atol=1e-9)
	import scipy.sparse.linalg as spla

	ref = spla.expm_multiply(A.tocsc(), V[:, 0])
	assert np.allclose(np.ravel(y), ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())


@pytest.mark.parametrize("orth,ncv", [(0, 2), (0, 20), (3, 20), (20, 20), (5, 5)])
def test_lanczos_dense_golden(eng, golden, orth, ncv):
	A, v = golden["A150"], golden["v42"]
	key = "lz_v42_deg20_orth5_nobasis" if (orth, ncv) == (5, 5) else f"lz_v42_deg20_orth{orth}"
	res = _run_lanczos(eng, A, v, 20, orth=orth, ncv=ncv, basis=(ncv == 20))
	assert rel_err(res["alpha"][0, :20], golden[key + "_a"]) < 1e-10
	assert rel_err(res["beta"][0, 1:20], golden[key + "_b"]) < 1e-10
	assert res["beta"][0, 0] == 0 and res["steps"][0] == 20 and abs(res["xnorm2"][0] - 150.0) < 1e-12
	if ncv == 20:
		Q = res["Q"][:, :, 0].T  # (n, ncv)
		assert np.allclose(Q, golden[key + "_Q"], atol=1e-9)


@pytest.mark.parametrize("orth", [0, 3])
def test_lanczos_block_matches_per_probe_golden(eng, golden, orth):
	"""a block of 4 probes == 4 independent reference runs (fresh basis per probe)."""
	A, V = golden["A150"], golden["V4_seed7"]
	res = _run_lanczos(eng, A, V, 20, orth=orth, ncv=20)
	assert rel_err(res["alpha"][:, :20], golden[f"quad_log_orth{orth}_alpha"][:, :20]) < 1e-10
	assert rel_err(res["beta"][:, 1:20], golden[f"quad_log_orth{orth}_beta"][:, 1:20]) < 1e-10


@pytest.mark.parametrize("orth", [0, 3])
def test_lanczos_sparse_golden(eng, golden, orth):
	L, V = _L30(golden), golden["VL_seed0"]
	Vp = np.concatenate([V, V[:, :1]], axis=1)  # pad to a power-of-two block
	res = _run_lanczos(eng, L, Vp, 50, orth=orth, ncv=50)
	assert rel_err(res["alpha"][:3, :50], golden[f"L30_quad_log_orth{orth}_alpha"][:, :50]) < 1e-10
	assert rel_err(res["beta"][:3, 1:50], golden[f"L30_quad_log_orth{orth}_beta"][:, 1:50]) < 1e-10
	assert np.array_equal(res["alpha"][3], res["alpha"][0])  # identical probes -> identical results (deterministic)


def test_lanczos_fp32_golden(eng, golden):
	L, V = _L30(golden), golden["VL_seed0"]
	res = _run_lanczos(eng, L.astype(np.float32), V[:, 0], 50, dtype=np.float32)
	assert res["alpha"].dtype == np.float32
	assert rel_err(res["alpha"][0, :50], golden["L30_lz32_a"]) < 1e-4
	assert rel_err(res["beta"][0, 1:50], golden["L30_lz32_b"]) < 1e-4
	res = _run_lanczos(eng, golden["A150"].astype(np.float32), golden["v42"], 20, dtype=np.float32)
	assert rel_err(res["alpha"][0, :20], golden["lz32_v42_deg20_a"]) < 1e-4
	assert rel_err(res["beta"][0, 1:20], golden["lz32_v42_deg20_b"]) < 1e-4


def test_lanczos_early_break_golden(eng, golden):
	"""lanczos.h:140: rank-3 operator -> break after 3 steps, alpha/beta tails stay zero."""
	B, V = golden["B60"], golden["VB_seed1"]
	res = _run_lanczos(eng, B, V, 10)
	for j in range(2):
		ga, gb = golden[f"B60_lz_alpha_{j}"], golden[f"B60_lz_beta_{j}"]
		assert res["steps"][j] == 3
		assert np.allclose(res["alpha"][j, :3], ga[:3], rtol=1e-10) and np.allclose(res["beta"][j, 1:3], gb[1:3], rtol=1e-10)
		assert np.all(res["alpha"][j, 3:] == 0) and np.all(res["beta"][j, 4:] == 0)
		assert res["beta"][j, 3] < np.sqrt(60) * 1e-8


def test_lanczos_vs_oracle_laplacian_blocks(eng, oracle):
	"""seeded Rademacher block on a 2-D Laplacian, k=32 device block vs the compiled reference per probe."""
	g = 64
	A = laplacian_2d(g)
	n = g * g
	st, inc = oracle.pcg64_state(11)
	X = eng.rademacher_block(st, inc, 0, 32, n, np.float64)
	Xh = X.cpu().numpy().copy()
	op = eng.DeviceOperator(A)
	out = op.lanczos(X, 50, 1e-8, 0, 2)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	W = oracle.ref_isotropic((n, 32), "rademacher", 11)
	assert np.array_equal(Xh, W)
	for p in (0, 1, 15, 31):
		ra, rb, _ = oracle.ref_lanczos(A, W[:, p], 50, 1e-8, 0)
		assert rel_err(al[p, :50], ra[:50]) < 1e-10 and rel_err(be[p, 1:50], rb[1:50]) < 1e-10


def test_quadrature_golden(eng, golden):
	import torch

	d, e = golden["T40_d"], golden["T40_e"]
	for deg, kn, kw in ((40, "T40_gw_nodes", "T40_gw_weights"), (25, "T40_gw25_nodes", "T40_gw25_weights")):
		a = torch.from_numpy(np.tile(d, (3, 1))).cuda()
		b = torch.from_numpy(np.tile(e, (3, 1))).cuda()
		theta, tau, y, info = eng.quadrature(a, b, deg, fun_id=0)
		theta, tau = theta.cpu().numpy(), tau.cpu().numpy()
		assert int(info.sum()) == 0
		assert np.allclose(theta[1], golden[kn], rtol=0, atol=1e-13)
		assert np.allclose(tau[1], golden[kw], rtol=0, atol=1e-12)
		assert abs(y.cpu().numpy()[2] - np.sum(golden[kn] * golden[kw])) < 1e-12
		assert abs(tau[0].sum() - 1.0) < 1e-13


@pytest.mark.parametrize("fun,kw", [("identity", {}), ("abs", {}), ("sqrt", {}), ("log", {}), ("inv", {}),
                                    ("exp", {"t": -0.5}), ("smoothstep", {"a": 0.2, "b": 0.6}),
                                    ("numrank", {"threshold": 0.3})])  # fmt: skip
@pytest.mark.parametrize("orth", [0, 3])
def test_quad_builtins_golden(eng, golden, fun, kw, orth):
	"""MatrixFunction.quad per probe (operators.py:126-151) == Lanczos block + fused quadrature."""
	A, V = golden["A150"], golden["V4_seed7"]
	op = eng.DeviceOperator(A)
	out = op.lanczos(eng.block_from_columns(V, np.float64), 20, 1e-8, orth, 20)
	fid, prm = eng.fun_spec(fun, kw)
	theta, tau, y, info = eng.quadrature(out["alpha"], out["beta"], 20, fid, prm, scale=out["xnorm2"])
	assert rel_err(y.cpu().numpy(), golden[f"quad_{fun}_orth{orth}"]) < 1e-8
	if fun == "log":
		assert np.allclose(theta.cpu().numpy(), golden[f"quad_log_orth{orth}_nodes"], atol=1e-11)
		assert np.allclose(tau.cpu().numpy(), golden[f"quad_log_orth{orth}_weights"], atol=1e-11)


def test_stats_update_matches_welford(eng, oracle):
	import torch

	rng = np.random.default_rng(0)
	y = rng.standard_normal(200) * 1e3 - 1.3e7
	w = oracle.Welford()
	state = torch.zeros(3, dtype=torch.float64, device="cuda")
	for lo in range(0, 200, 32):
		chunk = y[lo : lo + 32]
		w.update(chunk)
		eng.stats_update(torch.from_numpy(chunk).cuda(), state)
	s = state.cpu().numpy()
	assert s[0] == 200 and abs(s[1] / w.mu - 1) < 1e-14 and abs(s[2] / w.S - 1) < 1e-12
	assert abs(s[2] / 199 - np.var(y, ddof=1)) / np.var(y, ddof=1) < 1e-10


def test_callback_operator(eng, golden):
	from scipy.sparse.linalg import aslinearoperator

	A, v = golden["A150"], golden["v42"]
	res = _run_lanczos(eng, aslinearoperator(A), v, 20)
	assert rel_err(res["alpha"][0, :20], golden["lz_v42_deg20_orth0_a"]) < 1e-10


_REORTH_CASE = """
import sys, numpy as np
from tests.conftest import laplacian_2d
from primate_b200 import engine as eng
g, k, deg, dtype = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), np.dtype(sys.argv[4]).type
A = laplacian_2d(g)
X = eng.block_from_columns(np.random.default_rng(5).standard_normal((g * g, k)).astype(dtype), dtype)
out = eng.DeviceOperator(A, dtype=dtype).lanczos(X, deg, 1e-8, deg, deg + 1)
np.savez(sys.argv[5], alpha=out["alpha"].cpu().numpy(), beta=out["beta"].cpu().numpy())
"""


@pytest.mark.parametrize("g,k,dtype,tol", [(200, 64, "float64", 1e-10), (200, 8, "float64", 1e-10), (96, 1, "float64", 1e-10), (160, 64, "float32", 1e-4)])
def test_reorth_chain_matches_per_vector_launches(eng, oracle, tmp_path, g, k, dtype, tol):
	"""Full re-orthogonalisation (lanczos.h:52-66) as ONE cooperative kernel per step (v kept in shared memory, a grid
	barrier per visited vector) against the per-vector launch chain (SLQ_REORTH=launches, separate processes) and,
	for one probe, against the compiled reference."""
	import os
	import subprocess
	import sys

	root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
	res = {}
	for mode in ("chain", "launches"):
		path = str(tmp_path / f"{mode}.npz")
		env = dict(os.environ, SLQ_REORTH=mode)
		subprocess.run([sys.executable, "-c", _REORTH_CASE, str(g), str(k), "20", dtype, path], check=True, cwd=root, env=env, timeout=600)
		res[mode] = np.load(path)
	assert rel_err(res["chain"]["alpha"][:, :20], res["launches"]["alpha"][:, :20]) < tol
	assert rel_err(res["chain"]["beta"][:, 1:20], res["launches"]["beta"][:, 1:20]) < tol
	if dtype == "float64":
		A = laplacian_2d(g)
		V = np.random.default_rng(5).standard_normal((g * g, k))
		ra, rb, _ = oracle.ref_lanczos(A, V[:, k - 1], 20, 1e-8, 20)
		assert rel_err(res["chain"]["alpha"][k - 1, :20], ra[:20]) < tol and rel_err(res["chain"]["beta"][k - 1, 1:20], rb[1:20]) < tol
