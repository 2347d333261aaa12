"""Device normal / sphere probes (libslq: slq_probes_normal, slq_normals_scan, slq_probes_sphere_scale) against the
installed NumPy's Generator -- the arithmetic the reference delegates to (src/primate/random.py:32-41:
rng.standard_normal(out=W) on an F-ordered array; sphere = normals * sqrt(N) / ||column||).  Bit-exact, including the
ziggurat's wedge and tail branches, the data-dependent raw-draw count, random access to probe i, and the state the
user's Generator is left in."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _host(n, k, pdf, seed):
	from primate_b200.random import isotropic

	return isotropic((n, k), pdf=pdf, seed=seed)


def test_tables_are_the_installed_numpys():
	from primate_b200 import _ffi

	assert _ffi.lib().slq_normals_numpy_version().decode() == np.__version__, "regenerate: python scripts/make_ziggurat_tables.py"


def test_normal_block_bit_exact_8M_samples():
	"""n = 1e6 x 8 columns in one block: ~175 000 wedge samples and ~2 100 tail samples among them."""
	from primate_b200.random import ProbeStream

	n, k = 1_000_000, 8
	S = ProbeStream(n, "normal", seed=2024)
	X = S.block(k).cpu().numpy()
	W = np.empty((n, k), order="F")
	g = np.random.default_rng(2024)
	g.standard_normal(out=W)
	assert np.array_equal(X, W)
	assert np.abs(W).max() > 3.6541528853610088  # the tail branch was exercised
	# raw draws consumed == what the Generator consumed
	chk = np.random.default_rng(2024)
	chk.bit_generator.advance(S._raw_at[k])
	assert chk.bit_generator.state == g.bit_generator.state


@pytest.mark.parametrize("pdf", ["normal", "sphere"])
def test_random_access_and_dtypes(pdf):
	"""Blocks asked for out of order (rank sharding) equal the columns of one host draw; float32 is a cast of float64."""
	from primate_b200.random import ProbeStream

	n = 100_003
	W = _host(n, 45, pdf, 7)
	S = ProbeStream(n, pdf, seed=7)
	assert np.array_equal(S.block(4, 37).cpu().numpy(), W[:, 37:41])  # seek from the stream start
	assert np.array_equal(S.block(2, 41).cpu().numpy(), W[:, 41:43])  # continues from the remembered position
	assert np.array_equal(S.block(8, 3).cpu().numpy(), W[:, 3:11])
	assert np.array_equal(S.block(1, 0).cpu().numpy(), W[:, 0:1])
	S32 = ProbeStream(n, pdf, seed=7, dtype=np.float32)
	assert np.array_equal(S32.block(4, 16).cpu().numpy(), W[:, 16:20].astype(np.float32))


def test_seek_across_scan_chunks(monkeypatch):
	"""A far-away probe: the seek walks several scan chunks (chunk shrunk for the test) and lands on the exact draw."""
	from primate_b200 import random as prandom

	monkeypatch.setattr(prandom, "_SEEK_CHUNK", 1 << 16)
	n = 20_011
	W = _host(n, 40, "normal", 11)
	S = prandom.ProbeStream(n, "normal", seed=11)
	assert np.array_equal(S.block(2, 38).cpu().numpy(), W[:, 38:40])
	g = np.random.default_rng(11)
	g.standard_normal(40 * n)
	chk = np.random.default_rng(11)
	chk.bit_generator.advance(S._raw_at[40])
	assert chk.bit_generator.state == g.bit_generator.state


@pytest.mark.parametrize("n", [7, 129, 1000, 4097, 1_000_003])
def test_sphere_bit_exact(n):
	"""Column norms in NumPy's pairwise summation order: the scaled probes equal isotropic(pdf="sphere") bit for bit."""
	from primate_b200.random import ProbeStream

	k = 4
	W = _host(n, k, "sphere", 99)
	X = ProbeStream(n, "sphere", seed=99).block(k).cpu().numpy()
	assert np.array_equal(X, W)
	assert np.allclose(np.linalg.norm(X, axis=0), np.sqrt(n))


@pytest.mark.parametrize("pdf", ["normal", "sphere"])
def test_hutch_with_device_normals_leaves_generator_like_the_reference(golden, oracle, pdf):
	"""hutch(pdf=normal/sphere): estimate on the reference's probes, and the user's Generator advanced exactly as the
	reference's loop would leave it -- also when the adaptive criterion stops inside a speculated block."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import isotropic
	from primate_b200.trace import hutch

	A = golden["A150"]
	M = MatrixFunction(A, "log", deg=20, orth=0)
	g = np.random.default_rng(21)
	est = hutch(M, pdf=pdf, converge="count", count=40, seed=g)
	r = np.random.default_rng(21)
	W = isotropic((150, 40), pdf=pdf, seed=r)
	ref = oracle.ref_quad(A, W, "log", deg=20, orth=0).mean()
	assert abs(est / ref - 1) < 1e-8
	assert g.bit_generator.state == r.bit_generator.state
	# adaptive stop (per-probe cadence on speculative blocks): consumed < probes generated
	g2 = np.random.default_rng(5)
	est2, res = hutch(M, pdf=pdf, converge="confidence", confidence=0.9, atol=3.0, rtol=0.0, seed=g2, full=False), None
	r2 = np.random.default_rng(5)
	draw = isotropic(pdf=pdf, seed=r2)
	from primate_b200.estimators import ConfidenceCriterion, MeanEstimator

	est_ref, cc = MeanEstimator(covariance=True), ConfidenceCriterion(0.9, atol=3.0, rtol=0.0)
	while not cc(est_ref):
		est_ref.update(oracle.ref_quad(A, draw(size=(150, 1)), "log", deg=20, orth=0))
	assert abs(est2 / est_ref.estimate - 1) < 1e-8
	assert g2.bit_generator.state == r2.bit_generator.state
