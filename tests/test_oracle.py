"""Pins the CPU oracle: every layer (compiled reference ext, NumPy restatement, C restatement) against the
golden vectors produced by the unmodified reference package (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import rel_err


def _L30(golden):
    return sp.csr_matrix((golden["L30_data"], golden["L30_indices"], golden["L30_indptr"]), shape=(900, 900))


def test_pcg64_raw_matches_numpy(oracle, golden):
    st, inc = oracle.pcg64_state(0)
    assert np.array_equal(oracle.np_pcg64_raw(st, inc, 0, 64), golden["raw_seed0_first64"])
    assert np.array_equal(oracle.c_pcg64_raw(st, inc, 0, 64), golden["raw_seed0_first64"])
    # random access == PCG64.advance  
    bg = np.random.PCG64(0)
    bg.advance(10**12 + 7)
    assert np.array_equal(oracle.c_pcg64_raw(st, inc, 10**12 + 7, 16), bg.random_raw(16))
    assert np.array_equal(oracle.np_pcg64_raw(st, inc, 10**12 + 7, 16), oracle.c_pcg64_raw(st, inc, 10**12 + 7, 16))


@pytest.mark.parametrize("seed,shape,key", [(123, (1000, 5), "signs_seed123_1000x5"), (99, (257, 33), "signs_seed99_257x33")])
def test_rademacher_stream_layout(oracle, golden, seed, shape, key):
    """probe i == stream elements [i*N, (i+1)*N); sign == MSB of the raw output (random.py:22-29)."""
    N, k = shape
    W = oracle.ref_isotropic(shape, "rademacher", seed)
    assert np.array_equal((W > 0).astype(np.uint8), golden[key])
    st, inc = oracle.pcg64_state(seed)
    for i in (0, k // 2, k - 1):
        col = oracle.c_rademacher(st, inc, i * N, N)
        assert np.array_equal(col, W[:, i])
        raw = oracle.c_pcg64_raw(st, inc, i * N, N)
        assert np.array_equal((raw >> np.uint64(63)).astype(np.uint8), golden[key][:, i])


@pytest.mark.parametrize("orth", [0, 3, 20])
def test_lanczos_layers_vs_golden_dense(oracle, golden, orth):
    A, v = golden["A150"], golden["v42"]
    ga, gb, gQ = (golden[f"lz_v42_deg20_orth{orth}_{s}"] for s in "abQ")
    # compiled reference with a fresh full basis (return_basis=True => ncv = deg)
    al, be, Q = oracle.ref_lanczos(A, v, 20, 1e-8, orth, ncv=20)
    assert np.array_equal(al[:20], ga) and np.array_equal(be[1:20], gb) and np.array_equal(Q, gQ)
    for name, fn in (("np", oracle.np_lanczos), ("c", oracle.c_lanczos)):
        al, be, Q, steps = fn(A, v, 20, 1e-8, orth, ncv=20)
        assert steps == 20
        assert rel_err(al[:20], ga) < 1e-11, name
        assert rel_err(be[1:20], gb) < 1e-11, name


def test_lanczos_small_ncv_policy(oracle, golden):
    """lanczos.py:88-89: ncv = clip(orth, 2, deg) when the basis is not returned."""
    A, v = golden["A150"], golden["v42"]
    for fn in (oracle.ref_lanczos, oracle.np_lanczos, oracle.c_lanczos):
        out = fn(A, v, 20, 1e-8, 5)
        assert out[2].shape == (150, 5)
        assert rel_err(out[0][:20], golden["lz_v42_deg20_orth5_nobasis_a"]) < 1e-11
        assert rel_err(out[1][1:20], golden["lz_v42_deg20_orth5_nobasis_b"]) < 1e-11


def test_lanczos_sparse_and_fp32(oracle, golden):
    L = _L30(golden)
    V = golden["VL_seed0"]
    for orth in (0, 3):
        for j in range(3):
            for fn in (oracle.ref_lanczos, oracle.np_lanczos, oracle.c_lanczos):
                out = fn(L, V[:, j].copy(), 50, 1e-8, orth, ncv=50)
                assert rel_err(out[0], golden[f"L30_quad_log_orth{orth}_alpha"][j]) < 1e-10
                assert rel_err(out[1][1:], golden[f"L30_quad_log_orth{orth}_beta"][j][1:]) < 1e-10
    v32 = V[:, 0].astype(np.float32)
    for fn in (oracle.ref_lanczos, oracle.c_lanczos):
        out = fn(L.astype(np.float32), v32, 50, 1e-8, 0)
        assert out[0].dtype == np.float32
        assert rel_err(out[0][:50], golden["L30_lz32_a"]) < 1e-4
        assert rel_err(out[1][1:50], golden["L30_lz32_b"]) < 1e-4


def test_early_break(oracle, golden):
    """lanczos.h:140: stop once beta < sqrt(n)*rtol; the alpha/beta tails stay zero (fresh buffers)."""
    B, V = golden["B60"], golden["VB_seed1"]
    for j in range(2):
        ga, gb = golden[f"B60_lz_alpha_{j}"], golden[f"B60_lz_beta_{j}"]
        assert np.all(gb[4:] == 0) and np.all(ga[3:] == 0)
        for fn in (oracle.np_lanczos, oracle.c_lanczos):
            al, be, _, steps = fn(B, V[:, j].copy(), 10, 1e-8, 0)
            assert steps == 3
            assert np.allclose(al[:3], ga[:3], rtol=1e-12) and np.allclose(be[1:3], gb[1:3], rtol=1e-12)
            assert np.all(al[3:] == 0) and np.all(be[4:] == 0) and be[3] < np.sqrt(60) * 1e-8
    y = oracle.ref_quad(B, V, "log", deg=10, orth=0)
    assert rel_err(y, golden["B60_quad_log"]) < 1e-12
    y = oracle.ref_quad(B, V, "exp", deg=10, orth=0, lanczos=oracle.c_lanczos)
    assert rel_err(y, golden["B60_quad_exp"]) < 1e-10


def test_quadrature_and_tqli(oracle, golden):
    d, e = golden["T40_d"], golden["T40_e"]
    th, tau = oracle.ref_quadrature(d, e, 40)
    assert np.array_equal(th, golden["T40_gw_nodes"]) and np.array_equal(tau, golden["T40_gw_weights"])
    th, tau = oracle.ref_quadrature(d, e, 25)
    assert np.array_equal(th, golden["T40_gw25_nodes"]) and np.array_equal(tau, golden["T40_gw25_weights"])
    w, Z = oracle.np_tqli(d, e)
    assert np.allclose(w, golden["T40_tqli_w"], rtol=0, atol=1e-13)
    assert np.allclose(Z[0] ** 2, golden["T40_tqli_Z"][0] ** 2, rtol=0, atol=1e-13)
    # tqli (unsorted) agrees with stemr once sorted: tests/test_tridiagonal.py:26-44 in the reference
    o = np.argsort(w)
    assert np.allclose(w[o], golden["T40_eigh_w"], atol=1e-13)
    assert np.allclose((Z[0] ** 2)[o], golden["T40_eigh_Y0sq"], atol=1e-12)
    assert np.allclose(oracle.c_fttr(golden["T40_gw_nodes"], d, e, 40), golden["T40_fttr_direct"], rtol=1e-12)


@pytest.mark.parametrize("fun,kw", [("identity", {}), ("abs", {}), ("sqrt", {}), ("log", {}), ("inv", {}),
                                    ("exp", {"t": -0.5}), ("smoothstep", {"a": 0.2, "b": 0.6}),
                                    ("numrank", {"threshold": 0.3})])  # fmt: skip
def test_quad_builtins(oracle, golden, fun, kw):
    A, V = golden["A150"], golden["V4_seed7"]
    for orth in (0, 3):
        y = oracle.ref_quad(A, V, fun, deg=20, orth=orth, **kw)
        assert rel_err(y, golden[f"quad_{fun}_orth{orth}"]) < 1e-12
        y = oracle.ref_quad(A, V, fun, deg=20, orth=orth, lanczos=oracle.c_lanczos, **kw)
        assert rel_err(y, golden[f"quad_{fun}_orth{orth}"]) < 1e-9


def test_hutch_harness(oracle, golden):
    A = golden["A150"]
    est = oracle.ref_hutch(A, "log", deg=20, orth=0, count=200, seed=5)
    assert abs(est - golden["hutch_log_count200_seed5"].item()) < 1e-10
    est, samples = oracle.ref_hutch(A, "log", deg=20, orth=0, count=200, seed=5, batch=32, return_samples=True)
    assert len(samples) == 224
    assert np.allclose(samples, golden["hutch_log_count200_seed5_full_b32_samples"], rtol=1e-12)
    assert abs(est - golden["hutch_log_count200_seed5_full_b32"].item()) < 1e-10
    # sharded stream == one stream  (PCG64.advance(i*N))
    _, s2 = oracle.ref_hutch(A, "log", deg=20, orth=0, count=24, seed=5, batch=8, first_probe=200, return_samples=True)
    assert np.allclose(s2, golden["hutch_log_count200_seed5_full_b32_samples"][200:224], rtol=1e-12)
    L = _L30(golden)
    est = oracle.ref_hutch(L, "log", deg=50, orth=0, count=64, seed=0)
    assert abs(est / golden["L30_hutch_log_count64_seed0"].item() - 1) < 1e-12


def test_doc_known_answers(golden):
    """docs/guides/intro_trace.html:1072,1142."""
    assert golden["A150_trace"].item() == 74.88426089840914
    assert abs(golden["A150_logdet"].item() - (-151.5297634573417)) < 1e-12
    assert abs(golden["quad_log_v42"].item() - (-138.90995746812578)) < 1e-12


def test_block_driver_restatements_match_reference_goldens(oracle, golden):
	"""oracle.ref_hutchpp / ref_xdiag / ref_xtrace / ref_diag (restated host algebra of trace.py:119-315 and
	diagonal.py:11-138) against values produced by the unmodified reference on the same matrix and seeds."""
	A = golden["A150"]
	mm = lambda X: A @ X  # noqa: E731
	assert abs(oracle.ref_hutchpp(mm, 150, 30, seed=4)[0] - golden["hutchpp_A_m30_seed4"].item()) < 1e-10
	assert np.allclose(oracle.ref_xdiag(mm, 150, 20, seed=6), golden["xdiag_A_m20_seed6"], rtol=1e-12, atol=1e-13)
	assert abs(oracle.ref_xtrace(mm, 150, 10, seed=8)[0] - golden["xtrace_A_b10_seed8"].item()) < 1e-9
	assert np.allclose(oracle.ref_diag(mm, 150, 50, seed=10), golden["diag_A_seed10"], rtol=1e-12, atol=1e-13)


def test_ref_matvec_matches_reference_golden(oracle, golden):
	"""oracle.ref_matvec (operators.py:102-124 restated around the compiled extension) against the reference's _matvec."""
	A, v = golden["A150"], golden["v42"]
	for fun, kw in (("log", {}), ("exp", {"t": -0.5}), ("identity", {})):
		for orth in (0, 20):
			y = oracle.ref_matvec(A, v, fun, deg=20, orth=orth, **kw)
			g = golden[f"matvec_{fun}_orth{orth}_v42"]
			assert np.max(np.abs(y - np.ravel(g))) < 1e-12 * np.max(np.abs(g))


def test_ziggurat_restatement_and_tables_match_installed_numpy():
	"""oracle/zig_oracle.py (NumPy's random_standard_normal restated on the raw PCG64 stream, tables from
	tests/golden/ziggurat_tables.npz) reproduces Generator.standard_normal bit for bit, with the same number of raw draws;
	the committed tables are the installed NumPy's (the header compiled into libslq carries the same arrays)."""
	import os
	import re

	from oracle import zig_oracle

	ki, wi, fi = zig_oracle.tables()
	assert ki.shape == wi.shape == fi.shape == (256,) and ki[1] == 0 and fi[0] == 1.0
	for seed, n in ((0, 120_000), (99, 50_000)):
		got, used = zig_oracle.np_standard_normal(seed, n)
		g = np.random.default_rng(seed)
		ref = g.standard_normal(n)
		assert np.array_equal(got, ref)
		chk = np.random.default_rng(seed)
		chk.bit_generator.advance(used)
		assert chk.bit_generator.state == g.bit_generator.state and used > n
	root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
	hdr = open(os.path.join(root, "primate_b200", "csrc", "slq_ziggurat_tables.h")).read()
	words = [int(w, 16) for w in re.findall(r"0x([0-9A-F]{16})ull", hdr)]
	assert len(words) == 768
	assert np.array_equal(np.array(words[:256], dtype=np.uint64), ki)
	assert np.array_equal(np.array(words[256:512], dtype=np.uint64), wi.view(np.uint64))
	assert np.array_equal(np.array(words[512:], dtype=np.uint64), fi.view(np.uint64))


def test_pairwise_plan_reproduces_numpy_column_sums():
	"""primate_b200.random._pairwise_plan (the stack-machine form of NumPy's pairwise summation the sphere kernels replay)
	gives np.sum's bits for column sums of an F-ordered array (random.py:38: np.sum(out**2, axis=0))."""
	from primate_b200.random import _pairwise_plan

	def leaf(a):
		if len(a) < 8:
			res = 0.0
			for x in a:
				res += x
			return res
		r = list(a[:8])
		i = 8
		while i < len(a) - (len(a) % 8):
			for j in range(8):
				r[j] += a[i + j]
			i += 8
		res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
		for x in a[i:]:
			res += x
		return res

	rng = np.random.default_rng(3)
	for n in (5, 8, 129, 1000, 4097, 70_001):
		W = np.empty((n, 2), order="F")
		rng.standard_normal(out=W)
		sq = (W**2)[:, 1].tolist()
		starts, lens, red = _pairwise_plan(n)
		assert lens.sum() == n and red.sum() == len(starts) - 1
		stack = []
		for a, b, c in zip(starts, lens, red):
			stack.append(leaf(sq[a : a + b]))
			for _ in range(c):
				y, x = stack.pop(), stack.pop()
				stack.append(x + y)
		assert len(stack) == 1 and stack[0] == float(np.sum(W**2, axis=0)[1])
