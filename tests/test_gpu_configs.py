"""GPU parity of every BASELINE.json config at its OWN parameters (function, degree, orth, dtype, operator kind and the
kernel variants those select) against the reference's compiled extension (oracle/_ref) on identical probes.

  C1  dense SPD 1000^2, f = log, deg 20, 200 Rademacher probes, fp64            -- FULL size (SURVEY 8d generator)
  C2  2-D Laplacian 1000^2 grid, f = log, deg 50, 10 000 probes                  -- full size, a 64-probe stride of the job
  C3  R-MAT graph Laplacian, fp32, f = exp(-t x), deg 30                         -- scale 16 (n = 65 536), both row orders
  C4  dense Gram matrix, f = smoothstep, deg 100, orth 100 (full re-orth.)       -- 2048^2 of rank 512, k = 32
  C5  3-D 7-point Laplacian, f = inv, deg 80, quad / f(A)x / xdiag / hutch++     -- 48^3 (quad, matvec), 24^3 (block drivers)

Tolerances are the north-star's: alpha/beta 1e-10 relative in fp64 (1e-4 in fp32) over the first deg steps, estimates
1e-8 relative in fp64 (1e-4 in fp32).  The full-size C3/C4/C5 runs live in scripts/run_configs.py (property checks)."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import laplacian_2d, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _cuda():
	import torch

	assert torch.cuda.is_available(), "GPU tests need a CUDA device"


def _probes(n, k, seed, first=0):
	"""(device block, host columns) of Rademacher probes [first, first + k) of the stream `seed`."""
	from primate_b200.random import ProbeStream

	X = ProbeStream(n, "rademacher", seed=seed, dtype=np.float64).block(k, first)
	return X, X.cpu().numpy().copy()


# ---- C1 ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def c1():
	rng = np.random.default_rng(0)
	B = rng.standard_normal((1000, 1000))
	return B @ B.T / 1000 + 0.1 * np.eye(1000)


@pytest.mark.parametrize("orth", [0, 3])
def test_C1_dense_logdet_full_size(c1, oracle, orth):
	"""hutch log-det of the dense SPD matrix, all 200 probes: estimate vs the reference loop, per-probe alpha/beta."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	M = MatrixFunction(c1, "log", deg=20, orth=orth)
	est, res = hutch(M, converge="count", count=200, seed=1, full=True, batch=200, record=True)
	ref, samples = oracle.ref_hutch(c1, fun="log", deg=20, orth=orth, count=200, seed=1, return_samples=True)
	assert abs(est / ref - 1) < 1e-8
	assert rel_err(np.ravel(res.estimator.values), samples) < 1e-8
	assert abs(hutch(M, converge="count", count=200, seed=1) / ref - 1) < 1e-8  # the per-probe cadence path
	X, W = _probes(1000, 64, 1)
	_, out = M.quad_block(X, keep_tridiag=True)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	for p in (0, 31, 32, 63):
		ra, rb, _ = oracle.ref_lanczos(c1, W[:, p], 20, 1e-8, orth, ncv=20)
		assert rel_err(al[p, :20], ra[:20]) < 1e-10 and rel_err(be[p, 1:20], rb[1:20]) < 1e-10


# ---- C2 ---------------------------------------------------------------------------------------------------------------
@pytest.mark.slow
def test_C2_logdet_stride_of_the_10k_probe_job(oracle):
	"""The north-star job at full size on a 64-probe stride (every 157th probe of the 10 000): quadrature samples vs the
	reference extension to 1e-8; the full 10 000-probe gate (988 core-seconds on the CPU) is scripts/c2_full_parity.py."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	A = laplacian_2d(1000)
	n, deg = A.shape[0], 50
	st, inc = oracle.pcg64_state(0)
	M = MatrixFunction(A, "log", deg=deg, orth=0)
	idx = list(range(0, 10000, 157))[:64]
	got = np.array([M.quad_block(engine.rademacher_block(st, inc, i, 1, n, np.float64)).item() for i in idx])
	block = M.quad_block(engine.rademacher_block(st, inc, 9984, 16, n, np.float64)).cpu().numpy()  # the job's last block
	assert abs(block[15] / M.quad_block(engine.rademacher_block(st, inc, 9999, 1, n, np.float64)).item() - 1) < 1e-10
	for j in (0, 21, 42, 63):  # 1.2 s of reference time each
		sign = oracle.c_rademacher(st, inc, idx[j] * n, n)
		ry = oracle.ref_quad(A, sign, "log", deg=deg, orth=0)[0]
		assert abs(got[j] / ry - 1) < 1e-8
	assert np.all(np.isfinite(got)) and abs(got.mean() / 1.1669e6 - 1) < 5e-3  # log det of the 1000^2 Laplacian


# ---- C3 ---------------------------------------------------------------------------------------------------------------
def _rmat_laplacian_host(scale, edge_factor=16, seed=0, abcd=(0.57, 0.19, 0.19, 0.05)):
	"""SURVEY 8d's C3 generator on the host: R-MAT edges, symmetrised, self-loops and duplicates dropped, L = D - W (fp32)."""
	a, b, c, _ = abcd
	n, m = 1 << scale, edge_factor << scale
	rng = np.random.default_rng(seed)
	src, dst = np.zeros(m, dtype=np.int64), np.zeros(m, dtype=np.int64)
	for bit in range(scale):
		r = rng.random(m)
		src |= (r >= a + b).astype(np.int64) << bit
		dst |= (((r >= a) & (r < a + b)) | (r >= a + b + c)).astype(np.int64) << bit
	keep = src != dst
	W = sp.coo_matrix((np.ones(keep.sum(), dtype=np.float32), (src[keep], dst[keep])), shape=(n, n)).tocsr()
	W = ((W + W.T) > 0).astype(np.float32)
	L = (sp.diags(np.asarray(W.sum(axis=1)).ravel().astype(np.float32)) - W).tocsr().astype(np.float32)
	L.sort_indices()
	return L


@pytest.fixture(scope="module")
def c3():
	L = _rmat_laplacian_host(16)
	lam = float(sp.linalg.eigsh(L.astype(np.float64), k=1, which="LA", return_eigenvectors=False, tol=1e-4)[0])
	return L, 1.0 / lam


@pytest.mark.parametrize("reorder", [False, "balance"])
def test_C3_rmat_heat_trace_fp32(c3, oracle, reorder):
	"""fp32 heat-kernel quadrature on a power-law graph through the non-zero-balanced SpMM variant, in the caller's row
	order and in the row-balancing order: alpha/beta and samples vs the reference's fp32 path to 1e-4."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	L, t = c3
	n = L.shape[0]
	lens = np.diff(L.indptr)
	hub = int(np.argmax(lens))
	assert lens.max() > 1000  # hub rows far beyond the 64-non-zero lean path
	op = engine.DeviceOperator(L, dtype=np.float32, reorder=reorder)
	assert (op.perm is not None) == bool(reorder)
	M = MatrixFunction(op, "exp", deg=30, orth=0, dtype=np.float32, t=-t)
	_, W = _probes(n, 8, 3)
	W = W.astype(np.float32)
	W[:, 7] = 0.0
	W[hub, 7] = 1.0  # a start vector supported on the top hub: its Krylov space grows through the hub's 1000s of neighbours
	y, out = M.quad_block(engine.block_from_columns(W, np.float32), keep_tridiag=True)
	y, al, be = y.cpu().numpy(), out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	steps = out["steps"].cpu().numpy()
	L64 = L.astype(np.float64)
	for p in (0, 3, 6, 7):
		assert int(steps[p]) == 30
		# fp32 Lanczos on a power-law Laplacian (degrees 1 .. 9700) is rounding-chaotic: the reference's own fp32 run keeps
		# 4 digits of the exact coefficients for one or two steps and none after seven.  So the 1e-4 bar is
		# applied where it means something -- on the prefix where the reference's fp32 run is itself within 1e-4 of the
		# exact (fp64) coefficients of the same fp32-valued matrix -- and beyond it our run (dots accumulated in fp64) must
		# not be further from the exact coefficients than the reference's is.
		ra, rb, _ = oracle.ref_lanczos(L, W[:, p], 30, 1e-8, 0, ncv=2)
		ea, eb, _ = oracle.ref_lanczos(L64, W[:, p].astype(np.float64), 30, 1e-8, 0, ncv=2)
		for got, ref32, exact in ((al[p, :30], ra[:30], ea[:30]), (be[p, 1:30], rb[1:30], eb[1:30])):
			err_ref = np.abs(ref32 - exact) / np.abs(exact)
			err_got = np.abs(got - exact) / np.abs(exact)
			bad = np.nonzero(err_ref >= 1e-4)[0]
			J = int(bad[0]) if bad.size else len(exact)
			assert J >= 1 and np.all(err_got[:J] < 1e-4), (p, J, err_got[:J])
			assert np.all(np.abs(got[:J] - ref32[:J]) / np.abs(ref32[:J]) < 2e-4)
			worst_ref = np.maximum.accumulate(err_ref)
			assert np.all(err_got <= 10 * worst_ref + 1e-5), (p, err_got, err_ref)  # never worse than the reference's drift
		# the quadrature value is what the estimator consumes, and Gauss quadrature is stable under that drift
		ry = oracle.ref_quad(L, W[:, p], "exp", deg=30, orth=0, dtype=np.float32, t=-t)[0]
		ey = oracle.ref_quad(L64, W[:, p].astype(np.float64), "exp", deg=30, orth=0, t=-t)[0]
		assert abs(y[p] / ry - 1) < 1e-4 and abs(y[p] / ey - 1) < 1e-4


# ---- C4 ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def c4():
	rng = np.random.default_rng(0)
	X = rng.standard_normal((2048, 512))
	G = X @ X.T / 512
	lam = float(np.linalg.eigvalsh(G)[-1])
	return G, {"a": 1e-6 * lam, "b": 1e-3 * lam}


def test_C4_gram_numerical_rank_full_reorth(c4, oracle):
	"""smoothstep eigencount with deg = orth = 100 on a rank-512 Gram matrix, k = 32: the tiled dense kernel and the
	cooperative re-orthogonalisation chain vs the reference with a fresh basis (alpha/beta 1e-10, rank estimate 1e-8)."""
	from primate_b200.operators import MatrixFunction

	G, kw = c4
	n, deg = G.shape[0], 100
	M = MatrixFunction(G, "smoothstep", deg=deg, orth=deg, **kw)
	X, W = _probes(n, 32, 0)
	y, out = M.quad_block(X, keep_tridiag=True)
	y, al, be = y.cpu().numpy(), out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	assert int(out["steps"].cpu().numpy().min()) == deg
	ry = np.zeros(32)
	for p in range(32):
		v, extra = oracle.ref_quad(G, W[:, p], "smoothstep", deg=deg, orth=deg, return_all=True, **kw)
		ry[p] = v[0]
		if p in (0, 13, 31):
			# The eigenvalue 0 of this matrix has multiplicity 1536: in exact arithmetic the Krylov space holds one null
			# vector, in floating point rounding noise seeds further copies, and around the step where a copy emerges
			# (step ~48, then ~97) alpha/beta are determined by that noise -- the reference's compiled extension, its own
			# NumPy restatement (lanczos.py:211-238) and a C port disagree by 10-27% there and re-converge afterwards,
			# while the quadrature value agrees to 1e-15.  alpha/beta are therefore pinned on the prefix where
			# two CPU implementations of the reference algorithm still agree to 1e-12.
			ra, rb = extra[0][0], extra[0][1]
			na, nb, _, _ = oracle.np_lanczos(G, W[:, p], deg, 1e-8, deg, ncv=deg)
			dis = np.maximum(np.abs(na[:deg] - ra[:deg]) / np.abs(ra[:deg]), np.abs(nb[1 : deg + 1] - rb[1 : deg + 1]) / np.abs(rb[1 : deg + 1]))
			bad = np.nonzero(dis > 1e-12)[0]
			J = int(bad[0]) if bad.size else deg
			assert J >= 20, J
			assert rel_err(al[p, :J], ra[:J]) < 1e-10 and rel_err(be[p, 1:J], rb[1:J]) < 1e-10
	assert rel_err(y, ry) < 1e-8  # per-probe eigencount samples: insensitive to that noise
	assert abs(y.mean() / ry.mean() - 1) < 1e-8 and abs(ry.mean() - 512) < 40  # the numerical rank, within Monte-Carlo noise


# ---- C5 ---------------------------------------------------------------------------------------------------------------
def _laplacian_3d(g):
	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
	I = sp.identity(g)  # noqa: E741
	A = sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)
	A = A.tocsr().astype(np.float64)
	A.sort_indices()
	return A


def test_C5_laplacian3d_inverse_quad_and_matvec(oracle):
	"""f = inv, deg 80 on the 48^3 7-point Laplacian: quadratic forms (1e-8) and f(A) x through both engines -- basis-free
	two-pass and stored basis -- (1e-8 of ||y||) vs the reference's stored-basis formula (operators.py:102-151)."""
	from primate_b200.operators import MatrixFunction

	A = _laplacian_3d(48)
	n, deg = A.shape[0], 80
	M = MatrixFunction(A, "inv", deg=deg, orth=0)
	X, W = _probes(n, 16, 5)
	y = M.quad_block(X).cpu().numpy()
	for p in (0, 7, 15):
		ry = oracle.ref_quad(A, W[:, p], "inv", deg=deg, orth=0)[0]
		assert abs(y[p] / ry - 1) < 1e-8
	V = W[:, :3]
	Y = MatrixFunction(A, "inv", deg=deg, orth=0, basis="twopass") @ V
	Ys = MatrixFunction(A, "inv", deg=deg, orth=0, basis="stored") @ V
	for j in (0, 2):
		ref = oracle.ref_matvec(A, V[:, j], "inv", deg=deg, orth=0)
		assert np.max(np.abs(Y[:, j] - ref)) < 1e-8 * np.linalg.norm(ref)
		assert np.max(np.abs(Ys[:, j] - ref)) < 1e-8 * np.linalg.norm(ref)
	assert np.max(np.abs(np.ravel(M._matvec(V[:, 1])) - Y[:, 1])) < 1e-10 * np.linalg.norm(Y[:, 1])
	# A (A^-1 x) = x up to the Krylov truncation error of deg 80 on a grid of diameter 141
	resid = np.linalg.norm(A @ Y[:, 0] - V[:, 0]) / np.linalg.norm(V[:, 0])
	assert resid < 0.2


@pytest.mark.parametrize("pdf", ["rademacher", "sphere"])
def test_C5_xdiag_and_hutchpp_vs_reference_algebra(oracle, pdf):
	"""xdiag(M, 64) and hutchpp(M, 96) with every panel on the GPU vs the reference's host algebra (diagonal.py:99-138,
	trace.py:135-182) fed by the reference's own f(A) X (compiled extension + operators.py:102-124) on the same probes."""
	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutchpp

	A = _laplacian_3d(24)
	n, deg = A.shape[0], 80
	M = MatrixFunction(A, "inv", deg=deg, orth=0)
	mm = lambda X: oracle.ref_matvec(A, X, "inv", deg=deg, orth=0)  # noqa: E731
	d = xdiag(M, m=64, pdf=pdf, seed=2)
	rd = oracle.ref_xdiag(mm, n, 64, pdf=pdf, seed=2)
	assert np.max(np.abs(d - rd)) < 1e-8 * np.max(np.abs(rd))
	if pdf == "rademacher":
		est = hutchpp(M, m=96, mode="full", pdf=pdf, seed=4)
		rest, _ = oracle.ref_hutchpp(mm, n, 96, pdf=pdf, seed=4)
		assert abs(est / rest - 1) < 1e-8
