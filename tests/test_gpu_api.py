"""GPU parity of the reference-facing Python API (primate_b200.{lanczos,operators,trace,diagonal}) against golden
vectors produced by the unmodified reference, and -- at BASELINE.json's full C2 size -- against the oracle on sampled
probe indices plus size-independent properties.  Tolerances: alpha/beta 1e-10 rel (fp64), estimates 1e-8 rel."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import laplacian_2d, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _cuda():
	import torch

	assert torch.cuda.is_available(), "GPU tests need a CUDA device"


def _L30(golden):
	return sp.csr_matrix((golden["L30_data"], golden["L30_indices"], golden["L30_indptr"]), shape=(900, 900))


def test_lanczos_api_golden(golden):
	from primate_b200.lanczos import lanczos

	A = golden["A150"]
	a, b = lanczos(A, deg=10, seed=0, orth=0)
	assert rel_err(a, golden["lz_seed0_deg10_a"]) < 1e-10 and rel_err(b, golden["lz_seed0_deg10_b"]) < 1e-10
	for orth in (0, 3, 20):
		(a, b), Q = lanczos(A, v0=golden["v42"], deg=20, orth=orth, return_basis=True)
		assert rel_err(a, golden[f"lz_v42_deg20_orth{orth}_a"]) < 1e-10
		assert rel_err(b, golden[f"lz_v42_deg20_orth{orth}_b"]) < 1e-10
		assert Q.shape == (150, 20) and Q.flags["F_CONTIGUOUS"]
		assert np.max(np.abs(Q - golden[f"lz_v42_deg20_orth{orth}_Q"])) < 1e-9
	a, b = lanczos(A, v0=golden["v42"], deg=20, orth=5)
	assert rel_err(a, golden["lz_v42_deg20_orth5_nobasis_a"]) < 1e-10
	T = lanczos(A, v0=golden["v42"], deg=20, orth=0, sparse_mat=True)
	assert T.shape == (20, 20) and rel_err(T.diagonal(), golden["lz_v42_deg20_orth0_a"]) < 1e-10


def test_hutch_goldens_and_generator_state(golden):
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	A = golden["A150"]
	M = MatrixFunction(A, "log", deg=20, orth=0)
	rng = np.random.default_rng(5)
	est = hutch(M, converge="count", count=200, seed=rng)
	assert abs(est / golden["hutch_log_count200_seed5"].item() - 1) < 1e-8
	# the user's Generator is left where the reference leaves it: 200 probes x 150 raw draws
	chk = np.random.default_rng(5)
	chk.bit_generator.advance(200 * 150)
	assert rng.bit_generator.state["state"] == chk.bit_generator.state["state"]

	est, res = hutch(M, converge="count", count=200, seed=5, full=True, batch=32, record=True)
	assert abs(est / golden["hutch_log_count200_seed5_full_b32"].item() - 1) < 1e-8
	assert len(res.estimator) == 224
	assert rel_err(np.ravel(res.estimator.values), golden["hutch_log_count200_seed5_full_b32_samples"]) < 1e-8

	est, res = hutch(M, seed=11, full=True, batch=1, record=True)
	assert len(res.estimator) == int(golden["hutch_log_default_seed11_n"])
	assert abs(est / golden["hutch_log_default_seed11"].item() - 1) < 1e-8
	est = hutch(M, seed=11)  # per-probe cadence on speculative blocks stops at the same probe
	assert abs(est / golden["hutch_log_default_seed11"].item() - 1) < 1e-8

	est, res = hutch(M, converge="confidence", confidence=0.9, atol=2.0, rtol=0.0, seed=3, full=True, batch=8, record=True)
	assert len(res.estimator) == int(golden["hutch_log_conf_seed3_b8_n"])
	assert abs(est / golden["hutch_log_conf_seed3_b8"].item() - 1) < 1e-8

	assert abs(hutch(A, seed=2) / golden["hutch_A_default_seed2"].item() - 1) < 1e-8  # plain matrix: x^T A x on the device
	M3 = MatrixFunction(A, "exp", deg=12, orth=0, t=0.25)
	assert abs(hutch(M3, converge="count", count=64, seed=9) / golden["hutch_exp_count64_seed9"].item() - 1) < 1e-8
	ML = MatrixFunction(_L30(golden), "log", deg=50, orth=0)
	assert abs(hutch(ML, converge="count", count=64, seed=0) / golden["L30_hutch_log_count64_seed0"].item() - 1) < 1e-8


def test_hutch_host_callable_fun_and_normal_pdf(golden, oracle):
	"""User callables are evaluated on the host on the (k x deg) Ritz values; normal probes come from the host stream."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	A = golden["A150"]
	M = MatrixFunction(A, fun=lambda x: np.log(np.maximum(x, np.finfo(float).eps)), deg=20, orth=0)
	assert abs(hutch(M, converge="count", count=200, seed=5) / golden["hutch_log_count200_seed5"].item() - 1) < 1e-8
	Mn = MatrixFunction(A, "log", deg=20, orth=0)
	est = hutch(Mn, pdf="normal", converge="count", count=40, seed=21)
	W = np.empty((150, 40), order="F")
	np.random.default_rng(21).standard_normal(out=W)
	ref = oracle.ref_quad(A, W, "log", deg=20, orth=0).mean()
	assert abs(est / ref - 1) < 1e-8


def test_matrix_function_matvec_goldens(golden):
	from primate_b200.operators import MatrixFunction

	A, v = golden["A150"], golden["v42"]
	for fun, kw in (("log", {}), ("exp", {"t": -0.5}), ("identity", {})):
		for orth in (0, 20):
			M = MatrixFunction(A, fun, deg=20, orth=orth, **kw)
			y = np.ravel(M._matvec(v.copy()))
			g = golden[f"matvec_{fun}_orth{orth}_v42"]
			assert np.max(np.abs(y - g)) < 1e-8 * np.max(np.abs(g)), (fun, orth)
	M = MatrixFunction(A, "identity", deg=150, orth=150)
	V = np.random.default_rng(0).standard_normal((150, 5))
	assert np.allclose(M @ V, A @ V, atol=1e-8)  # SciPy LinearOperator interface
	assert np.allclose(M.quad(V), np.einsum("ij,ij->j", V, A @ V), rtol=1e-9)  # tests/test_operator.py:35-46


def test_block_drivers_on_matrix_function(golden):
	"""hutch++ / xdiag with a MatrixFunction: every `A @ block` is blocked device Lanczos (tests/test_trace.py:36-45)."""
	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutchpp

	A = golden["A150"]
	ew, U = np.linalg.eigh(A)
	expA = (U * np.exp(0.5 * ew)) @ U.T
	M = MatrixFunction(A, "exp", deg=40, orth=40, t=0.5)
	assert abs(hutchpp(M, m=60, seed=1) - np.trace(expA)) < 1e-2 * np.trace(expA)
	d = xdiag(M, m=120, seed=2, pdf="rademacher")
	assert np.linalg.norm(d - np.diag(expA)) < 0.1 * np.linalg.norm(np.diag(expA))


def test_hutchpp_device_panels_match_host_algebra(golden):
	"""hutch++ on a MatrixFunction keeps its panels on the GPU (device probes, device thin QR); the same call on the
	operator wrapped as a plain LinearOperator runs the reference's host algebra on the same probes.  Also the
	device-panel products themselves, and that the user's Generator ends where the host path leaves it."""
	import torch
	from scipy.sparse.linalg import LinearOperator

	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutchpp

	A = _L30(golden)
	n = A.shape[0]
	M = MatrixFunction(A, "exp", deg=30, orth=30, t=-0.3)
	X = np.random.default_rng(4).standard_normal((n, 37))
	Xd = torch.from_numpy(X).cuda()
	assert np.allclose(M.matmat_device(Xd).cpu().numpy(), M @ X, rtol=1e-11, atol=1e-11)
	assert np.allclose(M.quad_device(Xd).cpu().numpy(), M.quad(X), rtol=1e-11)
	assert np.array_equal(Xd.cpu().numpy(), X)  # panels are left intact
	for m in (30, 64):
		g1, g2 = np.random.default_rng(9), np.random.default_rng(9)
		dev, info = hutchpp(M, m=m, mode="full", seed=g1, full=True)
		plain = LinearOperator(M.shape, matvec=M._matvec, matmat=M._matmat, dtype=M.dtype)  # not a MatrixFunction: host algebra
		host, hinfo = hutchpp(plain, m=m, mode="full", seed=g2, full=True)
		assert abs(dev / host - 1) < 1e-9 and info.nit == hinfo.nit
		assert np.allclose(info.samples, hinfo.samples, rtol=1e-6, atol=1e-8 * abs(host))
		assert g1.integers(1 << 30) == g2.integers(1 << 30)
	red = hutchpp(M, m=64, seed=9)  # default mode: Lanczos quadrature for the sketched part
	assert abs(red / host - 1) < 1e-3


@pytest.mark.parametrize("pdf", ["rademacher", "sphere"])
def test_xdiag_device_panels_match_host_algebra(golden, pdf):
	"""xdiag on a MatrixFunction (panels on the GPU) against the reference's host algebra on the same probes."""
	from scipy.sparse.linalg import LinearOperator

	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction

	A = _L30(golden)
	M = MatrixFunction(A, "exp", deg=30, orth=30, t=-0.3)
	plain = LinearOperator(M.shape, matvec=M._matvec, matmat=M._matmat, rmatvec=M._matvec, rmatmat=M._matmat, dtype=M.dtype)
	for m in (20, 100):
		d_dev = xdiag(M, m=m, pdf=pdf, seed=11)
		d_host = xdiag(plain, m=m, pdf=pdf, seed=11)
		assert d_dev.shape == d_host.shape == (A.shape[0],)
		assert np.max(np.abs(d_dev - d_host)) < 1e-9 * np.max(np.abs(d_host))


@pytest.fixture(scope="module")
def c2():
	A = laplacian_2d(1000)
	return A


def test_c2_full_size_parity_on_sampled_probes(c2, oracle):
	"""n = 1e6, deg = 50: alpha/beta of probes at block and rank edges vs the reference extension (1e-10), their
	quadrature values (1e-8), and bit-exact signs."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction

	n, deg = c2.shape[0], 50
	st, inc = oracle.pcg64_state(0)
	M = MatrixFunction(c2, "log", deg=deg, orth=0)
	for first in (0, 4992, 9984):  # blocks holding probes 0, 31 | 4999 | 9999
		X = engine.rademacher_block(st, inc, first, 32, n, np.float64)
		Xh = X.cpu().numpy()
		y, out = M.quad_block(X, keep_tridiag=True)
		y, al, be = y.cpu().numpy(), out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
		for p in {0: (0, 31), 4992: (7,), 9984: (15,)}[first]:
			sign = oracle.c_rademacher(st, inc, (first + p) * n, n)
			assert np.array_equal(Xh[:, p], sign)
			ra, rb = oracle.ref_lanczos(c2, sign.copy(), deg, 1e-8, 0, ncv=2)[:2]
			assert rel_err(al[p, :deg], ra[:deg]) < 1e-10 and rel_err(be[p, 1:deg], rb[1:deg]) < 1e-10
			ry = oracle.ref_quad(c2, sign, "log", deg=deg, orth=0)[0]
			assert abs(y[p] / ry - 1) < 1e-8


def test_c2_full_size_properties(c2):
	"""Size-independent checks at n = 1e6: quad is invariant to the block a probe travels in and to its slot; tr(A) by
	identity-f hutch is exact for Rademacher probes on a matrix with constant diagonal up to the off-diagonal noise."""
	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	n = c2.shape[0]
	M = MatrixFunction(c2, "log", deg=50, orth=0)
	st, inc = 0x1234567890ABCDEF1234567890ABCDEF, (0xDEADBEEF << 1) | 1
	y32 = M.quad_block(engine.rademacher_block(st, inc, 64, 32, n, np.float64)).cpu().numpy()
	y8 = M.quad_block(engine.rademacher_block(st, inc, 72, 8, n, np.float64)).cpu().numpy()
	y1 = M.quad_block(engine.rademacher_block(st, inc, 95, 1, n, np.float64)).cpu().numpy()
	assert rel_err(y32[8:16], y8) < 1e-10 and abs(y32[31] / y1[0] - 1) < 1e-10
	Mi = MatrixFunction(c2, "identity", deg=2, orth=0)
	est = hutch(Mi, converge="count", count=64, seed=3)
	assert abs(est / 4e6 - 1) < 5e-3  # tr(A) = 4 n


def test_quad_range_equals_quad_block_and_quad(golden):
	"""The batched-quadrature range path (one Golub-Welsch launch for many Lanczos blocks) returns bit-identical samples
	to the per-block path and the reference-facing `quad` on the same probes."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import ProbeStream

	A = _L30(golden)
	M = MatrixFunction(A, "log", deg=20, orth=0)
	stream = ProbeStream(A.shape[0], "rademacher", seed=7, dtype=np.float64)
	y_range = M.quad_range(stream, 3, 77, 32).cpu().numpy()
	parts, first = [], 3
	for k in (32, 32, 8, 4, 1):
		parts.append(M.quad_block(stream.block(k, first)).cpu().numpy())
		first += k
	assert np.array_equal(y_range, np.concatenate(parts))
	W = np.concatenate([stream.block(k, f).cpu().numpy() for f, k in ((3, 32), (35, 32), (67, 8), (75, 4), (79, 1))], axis=1)
	assert rel_err(M.quad(W), y_range) < 1e-12


def test_device_csr_operator_matches_host_ingestion(golden):
	"""DeviceOperator.from_device_csr (torch CSR arrays already on the GPU; with and without the locality ordering, and with
	the row-length balancing order for power-law graphs) against the SciPy ingestion path: SpMM, hutch estimate and the host `matvec` protocol."""
	import torch

	from primate_b200 import engine
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	A = _L30(golden)
	n = A.shape[0]
	rp = torch.from_numpy(A.indptr.astype(np.int32)).cuda()
	ci = torch.from_numpy(A.indices.astype(np.int32)).cuda()
	va = torch.from_numpy(A.data.astype(np.float64)).cuda()
	ref = hutch(MatrixFunction(A, "exp", deg=20, orth=0), converge="count", count=64, seed=3)
	x = np.random.default_rng(0).standard_normal(n)
	for reorder in (False, 64, "balance"):
		op = engine.DeviceOperator.from_device_csr(rp, ci, va, n, dtype=np.float64, reorder=reorder)
		assert (op.perm is not None) == bool(reorder)
		assert np.allclose(op.matvec(x), A @ x, rtol=1e-13, atol=1e-13)
		est = hutch(MatrixFunction(op, "exp", deg=20, orth=0), converge="count", count=64, seed=3)
		assert abs(est / ref - 1) < 1e-10


def test_ingestion_formats_without_host_conversion(golden):
	"""CSC is taken as it is (a symmetric CSC matrix is its own CSR), COO and BSR are converted on the device: SpMM and
	Lanczos coefficients equal the CSR ingestion path; COO duplicates are summed as SciPy's tocsr() does."""
	from primate_b200 import engine

	A = _L30(golden)
	n = A.shape[0]
	X = np.random.default_rng(0).standard_normal((n, 8))
	ref = A @ X
	base = engine.DeviceOperator(A, reorder=False)
	o0 = base.lanczos(engine.block_from_columns(X, np.float64), 12, 1e-8, 0, 2)
	coo = A.tocoo()
	perm = np.random.default_rng(1).permutation(coo.nnz)
	dup = sp.coo_matrix((np.concatenate([coo.data[perm] * 0.25, coo.data[perm] * 0.75]), (np.tile(coo.row[perm], 2), np.tile(coo.col[perm], 2))), shape=A.shape)
	B = sp.kron(A[:300, :300], np.array([[2.0, 1.0, 0.0], [1.0, 2.0, 1.0], [0.0, 1.0, 2.0]])).tobsr(blocksize=(3, 3))
	for M in (A.tocsc(), coo, dup, A.tobsr(blocksize=(2, 2))):
		op = engine.DeviceOperator(M, reorder=False)
		assert op.nnz == A.nnz or M.format == "bsr"
		Y = op.matmat(engine.block_from_columns(X, np.float64)).cpu().numpy()
		assert np.max(np.abs(Y - ref)) < 1e-12 * np.max(np.abs(ref)), M.format
		o = op.lanczos(engine.block_from_columns(X, np.float64), 12, 1e-8, 0, 2)
		assert rel_err(o["alpha"].cpu().numpy()[:, :12], o0["alpha"].cpu().numpy()[:, :12]) < 1e-11
	opb = engine.DeviceOperator(B, reorder=False)
	Xb = np.random.default_rng(2).standard_normal((900, 4))
	assert np.max(np.abs(opb.matmat(engine.block_from_columns(Xb, np.float64)).cpu().numpy() - B @ Xb)) < 1e-11


def test_spectral_nodes_and_density(golden, oracle):
	"""Per-probe Gauss rules from the device (theta, tau for every probe of a range) equal Lanczos + LAPACK on the same
	probes; the smoothed density integrates to one and its moments match tr(A^j) / n."""
	from primate_b200.integrate import spectral_density
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import ProbeStream, isotropic

	A = golden["A150"]
	M = MatrixFunction(A, "identity", deg=20, orth=0)
	theta, tau, xn = M.nodes_range(ProbeStream(150, "rademacher", seed=4), 0, 40)
	theta, tau = theta.cpu().numpy(), tau.cpu().numpy()
	W = isotropic((150, 40), seed=4)
	for p in (0, 17, 39):
		a, b, _ = oracle.ref_lanczos(A, W[:, p], 20, 1e-8, 0)
		th, ta = oracle.ref_quadrature(a, b, 20)
		assert np.max(np.abs(theta[p] - th)) < 1e-12 and np.max(np.abs(tau[p] - ta)) < 1e-11
	assert np.allclose(xn.cpu().numpy(), 150.0)
	g = np.random.default_rng(4)
	t, dens, th2, ta2 = spectral_density(A, deg=20, count=40, points=400, seed=g, full=True)
	assert np.array_equal(th2, theta) and np.array_equal(ta2, tau)
	chk = np.random.default_rng(4)
	chk.bit_generator.advance(40 * 150)
	assert g.bit_generator.state == chk.bit_generator.state
	assert abs(np.trapezoid(dens, t) - 1.0) < 1e-3
	ew = np.linalg.eigvalsh(A)
	for j in (1, 2):  # moments of the quadrature measure: E[x^T A^j x] / n = tr(A^j) / n
		mom = float((tau * theta**j).sum(axis=1).mean())
		assert abs(mom - (ew**j).mean()) < 0.05 * (ew**j).mean()
